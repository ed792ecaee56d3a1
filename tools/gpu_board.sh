set -x
mkdir -p gpurun_out
timeout 600 python tools/bench_board.py > gpurun_out/bench_board.json 2> gpurun_out/bench_board.err; echo "board rc=$?"
timeout 600 ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_fmaheavy.sum,smsp__inst_executed_pipe_fmalite.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_lsu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,sm__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed -o gpurun_out/r1_board -f python tools/bench_board.py --ncu > gpurun_out/ncu_board.log 2>&1; echo "ncu rc=$?"
cat gpurun_out/bench_board.json; tail -5 gpurun_out/bench_board.err; tail -5 gpurun_out/ncu_board.log
