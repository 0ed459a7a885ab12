for v in 1 0 1 0; do
  JANUS_NVCC_EXTRA=-DGEMM_EPI_PREFETCH=$v python -m janus_b200.build --force > /dev/null 2>&1
  echo "epi prefetch $v: $(python scripts/gpu_dec_step_time.py | tail -1)"
done
