set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_stamps.txt
for lsu in 0,0 1,4; do
  echo "=== AZ_B200_FCSTACK_LSU=$lsu G=4096" >> gpurun_out/fc_stamps.txt
  timeout 240 env FCSTACK_STAMPS=1 AZ_B200_FCSTACK_LSU=$lsu FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps.txt
  echo "=== AZ_B200_FCSTACK_LSU=$lsu G=128" >> gpurun_out/fc_stamps.txt
  timeout 240 env FCSTACK_STAMPS=1 AZ_B200_FCSTACK_LSU=$lsu FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 128 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps.txt
done
cat gpurun_out/fc_stamps.txt
