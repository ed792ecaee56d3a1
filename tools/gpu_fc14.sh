set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_v3.txt
timeout 200 env FCSTACK_PROBE=v3 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_v3.txt
echo "=== stamps" >> gpurun_out/fc_v3.txt
timeout 200 env FCSTACK_STAMPS=1 AZ_B200_FCSTACK="2,4" FCSTACK_PROBE=v3 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" | head -30 >> gpurun_out/fc_v3.txt
cat gpurun_out/fc_v3.txt
timeout 600 python -m pytest tests/test_gpu_fusednet.py -x -q -k fcstack > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -15 gpurun_out/pytest_fused.log
