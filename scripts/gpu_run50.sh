timeout 600 python -m pytest tests/test_gpu_decoder.py tests/test_gpu_xlat.py -x -q 2>&1 | tail -2
echo "prefetch on : $(python scripts/gpu_dec_step_time.py | tail -1)"
JANUS_NVCC_EXTRA=-DJANUS_NO_L2_PREFETCH python -m janus_b200.build --force > /dev/null 2>&1
echo "prefetch off: $(python scripts/gpu_dec_step_time.py | tail -1)"
python -m janus_b200.build --force > /dev/null 2>&1
echo "prefetch on : $(python scripts/gpu_dec_step_time.py | tail -1)"
