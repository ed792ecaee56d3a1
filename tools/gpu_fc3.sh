set -x
mkdir -p gpurun_out
for mc in 0 1 2 3; do
  echo "=== AZ_B200_FCSTACK_MC=$mc" >> gpurun_out/fc_mc.txt
  timeout 240 env AZ_B200_FCSTACK_MC=$mc FCSTACK_PROBE=othello8 python tools/fcstack_probe.py 4096 >> gpurun_out/fc_mc.txt 2>&1
done
echo "=== small, MC=0" >> gpurun_out/fc_mc.txt
timeout 240 env AZ_B200_FCSTACK_MC=0 FCSTACK_PROBE=small python tools/fcstack_probe.py >> gpurun_out/fc_mc.txt 2>&1
cat gpurun_out/fc_mc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tools/net_microbench.py 4096 400 > gpurun_out/net_mb.txt 2>&1; sed -n 1,20p gpurun_out/net_mb.txt
