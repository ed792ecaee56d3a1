set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fusednet.py -x -q -k "trunk2" > gpurun_out/pytest_t3.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t3.log
tail -15 gpurun_out/pytest_t3.log
timeout 300 python tools/trunk2_probe.py 4096 3 > gpurun_out/trunk3_probe.txt 2>&1; cat gpurun_out/trunk3_probe.txt
timeout 600 env AZ_B200_TRUNK_V3=1 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 > gpurun_out/bench_t3.json 2> gpurun_out/bench_t3.err; tail -c 700 gpurun_out/bench_t3.json
