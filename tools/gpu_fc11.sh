set -x
mkdir -p gpurun_out
rm -f gpurun_out/fc_stamps5.txt
echo "=== two MMA warps, G=4096" >> gpurun_out/fc_stamps5.txt
timeout 240 env FCSTACK_STAMPS=1 FCSTACK_PROBE=othello8_14 python tools/fcstack_probe.py 4096 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps5.txt
echo "=== small default" >> gpurun_out/fc_stamps5.txt
timeout 240 env FCSTACK_PROBE=small python tools/fcstack_probe.py 2>&1 | grep -v "Warning\|capture_end" >> gpurun_out/fc_stamps5.txt
cat gpurun_out/fc_stamps5.txt
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py tests/test_gpu_parity_evidence.py::test_othello_net_matches_reference_outputs -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -5 gpurun_out/pytest_fused.log
timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_quick.json; python -c "
import json; d=json.loads(open('gpurun_out/bench_quick.json').read().strip().split(chr(10))[-1]); print(d['value']/1e6, d['roofline']['kernel_ms_alone_in_graph'], d['roofline']['ms_per_launch'], d['roofline']['ms_per_search_step'], d['roofline_tree']['ms_per_launch'])"
