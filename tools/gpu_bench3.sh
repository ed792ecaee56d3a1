set -x
mkdir -p gpurun_out
for i in 1 2 3 4; do
timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 4 --warmup 3 > gpurun_out/bench_rep$i.json 2> gpurun_out/bench_rep$i.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_rep$i.json').read().strip().split('\n')[-1]); r=d['roofline']; print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), r['per_step_ms'], round(r['search_share_of_batch'],4))"
done
