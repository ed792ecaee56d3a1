# round-2 ncu evidence: (1) launch list of the product path, (2) --set full of the three dominant kernels.
# Each ncu command line is first run WITHOUT ncu (B200_PROFILING.md).
set -x
mkdir -p gpurun_out
CMD="python bench.py --games 4096 --sims 8 --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_net_trunk|k_net_fcstack|k_tree_step" -s 600 -c 6 -o gpurun_out/r2_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
