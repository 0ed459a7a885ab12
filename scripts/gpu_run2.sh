mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q 2>&1 | tail -25
timeout 300 python scripts/gpu_gemm_bench.py 2>&1 | tail -12
