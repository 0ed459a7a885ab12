JANUS_XLAT_HALF=1 timeout 600 python -m pytest tests/test_gpu_xlat.py -x -q 2>&1 | tail -4
for v in 0 1 0 1; do echo "half $v: $(JANUS_XLAT_HALF=$v timeout 120 python scripts/gpu_xlat_bench.py | head -1)"; done
JANUS_XLAT_HALF=1 timeout 600 python -m pytest tests/test_gpu_decoder.py -x -q -k "latent" 2>&1 | tail -2
