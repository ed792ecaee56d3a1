set -x
mkdir -p gpurun_out
CMD="python bench.py --games 4096 --sims 8 --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
export AZ_B200_TRUNK_V2=1
timeout 600 $CMD > gpurun_out/ncu_t2_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_t2_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_net_trunk2" -s 300 -c 2 -o gpurun_out/r2_trunk2 -f $CMD > gpurun_out/ncu_t2.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_t2.log
ls -la gpurun_out/r2_trunk2.ncu-rep
