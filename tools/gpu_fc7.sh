set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_stamps3.txt
for mc in 0 1; do
  echo "=== AZ_B200_FCSTACK_MC=$mc G=4096" >> gpurun_out/fc_stamps3.txt
  timeout 240 env FCSTACK_STAMPS=1 AZ_B200_FCSTACK_MC=$mc FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps3.txt
done
echo "=== small default" >> gpurun_out/fc_stamps3.txt
timeout 240 env FCSTACK_PROBE=small python tools/fcstack_probe.py 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps3.txt
cat gpurun_out/fc_stamps3.txt
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -5 gpurun_out/pytest_fused.log
timeout 1500 python -m pytest tests/test_gpu_parity_evidence.py -x -q > gpurun_out/pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_parity.log
tail -25 gpurun_out/pytest_parity.log
