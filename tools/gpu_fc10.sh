set -x
mkdir -p gpurun_out
timeout 300 python tools/net_microbench.py 4096 400 > gpurun_out/net_mb.txt 2>&1; sed -n 1,24p gpurun_out/net_mb.txt
timeout 300 env AZ_B200_NO_PDL=1 python tools/net_microbench.py 4096 400 > gpurun_out/net_mb_nopdl.txt 2>&1; sed -n 1,24p gpurun_out/net_mb_nopdl.txt
