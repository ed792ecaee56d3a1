set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -5 gpurun_out/pytest_noalloc_alone.log
timeout 900 python -m pytest tests/test_gpu_fusednet.py tests/test_gpu_connect4.py tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -5 gpurun_out/pytest_fused.log
