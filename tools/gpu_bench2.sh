set -x
mkdir -p gpurun_out
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cat gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
timeout 900 python bench.py --no-cpu-baseline --sims 400 --steps 1 --warmup 3 > gpurun_out/bench_sims400.json 2> gpurun_out/bench_sims400.err; echo "bench400 rc=$?"; cat gpurun_out/bench_sims400.json; tail -5 gpurun_out/bench_sims400.err
