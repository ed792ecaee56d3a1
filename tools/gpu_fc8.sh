set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_stamps4.txt
echo "=== two producer warps, G=4096" >> gpurun_out/fc_stamps4.txt
timeout 240 env FCSTACK_STAMPS=1 FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps4.txt
echo "=== small default" >> gpurun_out/fc_stamps4.txt
timeout 240 env FCSTACK_PROBE=small python tools/fcstack_probe.py 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps4.txt
cat gpurun_out/fc_stamps4.txt
timeout 1800 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_parity_evidence.py > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cat gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
