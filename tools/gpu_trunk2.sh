set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fusednet.py -x -q -k "trunk2 or large_batch or trunk_layers" > gpurun_out/pytest_trunk2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_trunk2.log
tail -25 gpurun_out/pytest_trunk2.log
