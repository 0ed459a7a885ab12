mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; tail -5 gpurun_out/gpu_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_stt_a.log 2>&1; tail -3 gpurun_out/bench_stt_a.log | cut -c1-3000
