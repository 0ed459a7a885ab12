mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_stt.py tests/test_gpu_decoder.py tests/test_gpu_service.py -q -x > gpurun_out/gpu_tests.log 2>&1; tail -4 gpurun_out/gpu_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_stt_d.log 2>&1; tail -1 gpurun_out/bench_stt_d.log | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['windows_per_gpu'], d['value'], d['e2e']['value'], d['stages_ms'], d['roofline']['frac'], d['roofline']['encoder_tensor']['frac'])"
