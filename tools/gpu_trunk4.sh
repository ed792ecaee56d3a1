set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -5 gpurun_out/pytest_fused.log
timeout 300 python tools/trunk2_probe.py > gpurun_out/trunk2_probe.txt 2>&1; cat gpurun_out/trunk2_probe.txt
timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 > gpurun_out/bench_t1.json 2> gpurun_out/bench_t1.err; tail -c 600 gpurun_out/bench_t1.json
timeout 600 env AZ_B200_TRUNK_V2=1 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 > gpurun_out/bench_t2.json 2> gpurun_out/bench_t2.err; tail -c 600 gpurun_out/bench_t2.json
