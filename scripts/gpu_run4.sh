mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decoder.py -q -s > gpurun_out/dec_test.log 2>&1
grep -E "graph=|logit err|min oracle|passed|failed|Error|error" gpurun_out/dec_test.log | head -30
