set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -6 gpurun_out/pytest_fused.log
timeout 300 python tools/net_microbench.py > gpurun_out/net_mb.txt 2>&1; tail -25 gpurun_out/net_mb.txt
timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 > gpurun_out/bench_t1.json 2> gpurun_out/bench_t1.err; tail -c 1200 gpurun_out/bench_t1.json
