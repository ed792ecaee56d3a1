set -x
mkdir -p gpurun_out
N=8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --sims 400 --steps 1 --warmup 3 > gpurun_out/bench_config4_n8.json 2> gpurun_out/bench_config4_n8.err; echo "config4 rc=$?"
grep '^{' gpurun_out/bench_config4_n8.json | cut -c1-300; tail -3 gpurun_out/bench_config4_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --game connect4 --games 2048 --steps 2 --warmup 3 > gpurun_out/bench_config5_n8.json 2> gpurun_out/bench_config5_n8.err; echo "config5 rc=$?"
grep '^{' gpurun_out/bench_config5_n8.json | cut -c1-300; tail -3 gpurun_out/bench_config5_n8.err
