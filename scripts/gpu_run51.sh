timeout 900 python -m pytest tests/test_gpu_decoder.py tests/test_gpu_gemm.py tests/test_gpu_service.py -x -q 2>&1 | tail -2
for v in 1 0 1 0; do echo "pdl $v: $(PDL=$v python scripts/gpu_dec_step_time.py | tail -1)"; done
