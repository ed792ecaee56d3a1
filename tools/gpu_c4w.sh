set -x
mkdir -p gpurun_out
rm -f gpurun_out/c4w.txt
for w in 4 28; do
  echo "=== connect4 AZ_B200_TREE_WARPS=$w" >> gpurun_out/c4w.txt
  timeout 600 env AZ_B200_TREE_WARPS=$w python bench.py --game connect4 --no-cpu-baseline --no-extras --steps 3 --warmup 4 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]); r=d['roofline']; print(round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'fwd', round(r['ms_per_launch']*1e3,2), 'step', round(r['ms_per_search_step']*1e3,2), 'tree', round(d['roofline_tree']['ms_per_launch']*1e3,2), 'share', round(r['search_share_of_batch'],4))" >> gpurun_out/c4w.txt
done
cat gpurun_out/c4w.txt
