set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_selfplay.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
timeout 900 python bench.py --no-cpu-baseline --game connect4 > gpurun_out/bench_quick_c4.json 2> gpurun_out/bench_quick_c4.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_quick_c4.json; tail -5 gpurun_out/bench_quick_c4.err
