mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r38.log 2>&1; tail -1 gpurun_out/bench_r38.log | cut -c1-1800
for m in 1 0 2 1; do
  JANUS_NVCC_EXTRA=-DATT_POLY_MODE=$m python -m janus_b200.build --force > /dev/null 2>&1
  echo "poly mode $m: $(timeout 300 python scripts/gpu_att_profile.py 2>&1 | grep kernel)"
done
