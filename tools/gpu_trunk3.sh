set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fusednet.py -x -q -k "trunk2 or large_batch or trunk_layers or within_1e2" > gpurun_out/pytest_trunk2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_trunk2.log
tail -6 gpurun_out/pytest_trunk2.log
timeout 300 python tools/trunk2_probe.py > gpurun_out/trunk2_probe.txt 2>&1; cat gpurun_out/trunk2_probe.txt
