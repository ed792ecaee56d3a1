set -x
mkdir -p gpurun_out
timeout 300 python tools/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
cat gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool ${1:-memcheck} --error-exitcode 9 python tools/sanitize_target.py > gpurun_out/sanitize_${1:-memcheck}.log 2>&1; echo "sanitizer rc=$?"
tail -15 gpurun_out/sanitize_${1:-memcheck}.log
