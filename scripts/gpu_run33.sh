mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stt.py -x -q 2>&1 | tail -5
for m in 2 0 3 2; do
  JANUS_NVCC_EXTRA=-DATT_POLY_MODE=$m python -m janus_b200.build --force > /dev/null 2>&1
  echo "poly mode $m: $(timeout 300 python scripts/gpu_gemm_bench.py 2>&1 | grep -i att)"
done
JANUS_NVCC_EXTRA=-DATT_PROFILE python -m janus_b200.build --force 2>&1 | tail -1
timeout 300 python scripts/gpu_att_profile.py 2>&1 | tail -9
