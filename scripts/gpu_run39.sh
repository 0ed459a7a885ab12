timeout 600 python -m pytest tests/test_gpu_xlat.py -x -q 2>&1 | tail -15
