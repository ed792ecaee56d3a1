set -x
mkdir -p gpurun_out
rm -f gpurun_out/pdl_ab.txt
for cfg in "0 0" "1 0" "0 1" "1 1"; do
  set -- $cfg
  echo "=== AZ_B200_FCSTACK_LATE_TRIGGER=$1 AZ_B200_TRUNK_LATE_TRIGGER=$2" >> gpurun_out/pdl_ab.txt
  timeout 600 env AZ_B200_FCSTACK_LATE_TRIGGER=$1 AZ_B200_TRUNK_LATE_TRIGGER=$2 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]); r=d['roofline']; print(round(d['value']/1e6,2), 'fwd', round(r['ms_per_launch']*1e3,2), 'step', round(r['ms_per_search_step']*1e3,2), 'tree', round(d['roofline_tree']['ms_per_launch']*1e3,2))" >> gpurun_out/pdl_ab.txt
done
echo "=== NO_PDL" >> gpurun_out/pdl_ab.txt
timeout 600 env AZ_B200_NO_PDL=1 python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]); r=d['roofline']; print(round(d['value']/1e6,2), 'fwd', round(r['ms_per_launch']*1e3,2), 'step', round(r['ms_per_search_step']*1e3,2), 'tree', round(d['roofline_tree']['ms_per_launch']*1e3,2))" >> gpurun_out/pdl_ab.txt
cat gpurun_out/pdl_ab.txt
