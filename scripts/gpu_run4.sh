mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decoder.py -x -q -s 2>&1 | tail -40
