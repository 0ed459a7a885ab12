mkdir -p gpurun_out
timeout 600 python scripts/gpu_dec_bench.py 2>&1 | tail -16
