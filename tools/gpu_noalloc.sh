set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -5 gpurun_out/pytest_noalloc_alone.log
timeout 3000 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
