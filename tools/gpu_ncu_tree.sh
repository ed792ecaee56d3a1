set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_tree_step" --launch-skip 8000 --launch-count 3 -o gpurun_out/r1e_tree -f python bench.py --games 4096 --sims 200 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_tree.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_tree.log
