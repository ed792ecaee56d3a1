set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_lsu.txt
for lsu in 0,0 1,0 1,2 1,4 1,6 1,8 0,8 0,4; do
  echo "=== AZ_B200_FCSTACK_LSU=$lsu" >> gpurun_out/fc_lsu.txt
  timeout 240 env AZ_B200_FCSTACK_LSU=$lsu FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_lsu.txt
done
echo "=== small default" >> gpurun_out/fc_lsu.txt
timeout 240 env FCSTACK_PROBE=small python tools/fcstack_probe.py 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_lsu.txt
cat gpurun_out/fc_lsu.txt
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -5 gpurun_out/pytest_fused.log
timeout 1200 python -m pytest tests/test_gpu_tree.py tests/test_gpu_throughput.py -x -q > gpurun_out/pytest_tp.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tp.log
tail -30 gpurun_out/pytest_tp.log
