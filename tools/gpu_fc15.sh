set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_v3.txt
timeout 200 env FCSTACK_STAMPS=1 FCSTACK_PROBE=v3 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" | head -32 >> gpurun_out/fc_v3.txt
cat gpurun_out/fc_v3.txt
