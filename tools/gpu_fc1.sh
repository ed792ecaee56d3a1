set -x
mkdir -p gpurun_out
timeout 240 env FCSTACK_PROBE=small python tools/fcstack_probe.py > gpurun_out/fc_small.txt 2>&1; echo "rc=$?" >> gpurun_out/fc_small.txt
tail -30 gpurun_out/fc_small.txt
timeout 240 env FCSTACK_PROBE=othello8 python tools/fcstack_probe.py 4096 > gpurun_out/fc_o8.txt 2>&1; echo "rc=$?" >> gpurun_out/fc_o8.txt
tail -12 gpurun_out/fc_o8.txt
timeout 900 python -m pytest tests/test_gpu_fusednet.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -15 gpurun_out/pytest_fused.log
