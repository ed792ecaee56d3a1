mkdir -p gpurun_out
for c in 8192 12288 16384 24576 38592; do
  timeout 200 python bench.py --no-cpu-baseline --steps 2 --warmup 2 --capacity $c > gpurun_out/cap_$c.json 2> gpurun_out/cap_$c.err
  echo "cap $c rc=$?"; tail -2 gpurun_out/cap_$c.err | cut -c1-300
  python - <<P
import json
d=json.loads(open('gpurun_out/cap_$c.json').read().strip().splitlines()[-1])
print($c, round(d['value']/1e6,2), 'step', d['roofline']['ms_per_search_step'], 'net', d['roofline']['ms_per_launch'], 'tree', d['roofline_tree']['ms_per_launch'], d['config']['node_pool'])
P
done
