for c in 3072 4096 6144 9648; do
  timeout 200 python bench.py --no-cpu-baseline --steps 1 --warmup 3 --capacity $c > gpurun_out/cap_$c.json 2> gpurun_out/cap_$c.err
  echo "cap $c rc=$?"; tail -2 gpurun_out/cap_$c.err | cut -c1-300; cut -c1-120 gpurun_out/cap_$c.json
done
