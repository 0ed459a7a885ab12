mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stt.py tests/test_gpu_gemm.py -x -q 2>&1 | tail -3
timeout 300 python scripts/gpu_gemm_bench.py 2>&1 | tail -8
JANUS_NVCC_EXTRA="-DATT_PROFILE -DATT_POLY_MODE=3" python -m janus_b200.build --force 2>&1 | tail -1
timeout 300 python scripts/gpu_att_profile.py 2>&1 | tail -11
