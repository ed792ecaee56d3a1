set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_stamps2.txt
for lsu in 1,0 0,0; do
  echo "=== AZ_B200_FCSTACK_LSU=$lsu G=4096" >> gpurun_out/fc_stamps2.txt
  timeout 240 env FCSTACK_STAMPS=1 AZ_B200_FCSTACK_LSU=$lsu FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps2.txt
done
echo "=== small default" >> gpurun_out/fc_stamps2.txt
timeout 240 env FCSTACK_PROBE=small python tools/fcstack_probe.py 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps2.txt
cat gpurun_out/fc_stamps2.txt
timeout 900 python -m pytest tests/test_gpu_fusednet.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -5 gpurun_out/pytest_fused.log
