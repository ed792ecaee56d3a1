set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n1.json
timeout 900 env AZ_B200_HEADS=library python bench.py --no-cpu-baseline --leaves 4 > gpurun_out/bench_n1_library_heads.json 2> gpurun_out/bench_n1_library_heads.err; echo "bench lib rc=$?"; cut -c1-300 gpurun_out/bench_n1_library_heads.json
timeout 600 python bench.py --game connect4 --no-cpu-baseline --no-extras > gpurun_out/bench_connect4.json 2> gpurun_out/bench_connect4.err; echo "c4 rc=$?"; cut -c1-300 gpurun_out/bench_connect4.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_reference.json
