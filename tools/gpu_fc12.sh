set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_v2.txt
timeout 300 env FCSTACK_PROBE=v2 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_v2.txt
echo "=== stamps" >> gpurun_out/fc_v2.txt
timeout 300 env FCSTACK_STAMPS=1 FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_v2.txt
echo "=== first kernel (AZ_B200_FCSTACK_V1=1)" >> gpurun_out/fc_v2.txt
timeout 300 env AZ_B200_FCSTACK_V1=1 FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_v2.txt
cat gpurun_out/fc_v2.txt
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -15 gpurun_out/pytest_fused.log
timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 > gpurun_out/bench_v2.json 2> gpurun_out/bench_v2.err; tail -c 1500 gpurun_out/bench_v2.json
