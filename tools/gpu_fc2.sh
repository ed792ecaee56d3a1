set -x
mkdir -p gpurun_out
timeout 240 env FCSTACK_PROBE=othello8 python tools/fcstack_probe.py 4096 > gpurun_out/fc_o8.txt 2>&1; echo "rc=$?" >> gpurun_out/fc_o8.txt
tail -8 gpurun_out/fc_o8.txt
timeout 240 python tools/fcstack_probe.py 16384 > gpurun_out/fc_o8_16k.txt 2>&1; tail -14 gpurun_out/fc_o8_16k.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tools/net_microbench.py 4096 400 > gpurun_out/net_mb.txt 2>&1; tail -30 gpurun_out/net_mb.txt
