set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_selfplay.py tests/test_gpu_dropin.py -m gpu -x -q 2>&1 | tail -15
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cat gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
