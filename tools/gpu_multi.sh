set -x
mkdir -p gpurun_out
N=${1:-2}
for game in othello connect4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/dist_train_check.py $game > gpurun_out/dist_train_$game.log 2>&1; echo "dist $game rc=$?"; tail -3 gpurun_out/dist_train_$game.log
done
[ -n "$SKIP_BENCH" ] || timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
cat gpurun_out/bench_n$N.json | cut -c1-600
