set -x
mkdir -p gpurun_out
rm -f gpurun_out/treew.txt
for w in 4 8 14 28; do
  echo "=== AZ_B200_TREE_WARPS=$w" >> gpurun_out/treew.txt
  timeout 600 env AZ_B200_TREE_WARPS=$w python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]); r=d['roofline']; print(round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'fwd', round(r['ms_per_launch']*1e3,2), 'step', round(r['ms_per_search_step']*1e3,2), 'tree', round(d['roofline_tree']['ms_per_launch']*1e3,2))" >> gpurun_out/treew.txt
done
cat gpurun_out/treew.txt
timeout 900 python -m pytest tests/test_gpu_tree.py -x -q 2>&1 | tail -3
