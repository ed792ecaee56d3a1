set -x
mkdir -p gpurun_out
for i in 1 2 3; do
timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 3 --warmup 3 > gpurun_out/bench_rep$i.json 2> gpurun_out/bench_rep$i.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_rep$i.json').read().strip().split('\n')[-1]); r=d['roofline']; print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), round(r['ms_per_launch']*1e3,2), round(r['ms_per_search_step']*1e3,2), r.get('kernel_ms_alone_in_graph'), round(d['roofline_tree']['ms_per_launch']*1e3,2))"
done
