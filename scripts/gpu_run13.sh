mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_stt_l64.log 2>&1; tail -3 gpurun_out/bench_stt_l64.log | cut -c1-2500
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lanes 4 > gpurun_out/bench_stt_l64x4.log 2>&1; tail -1 gpurun_out/bench_stt_l64x4.log | cut -c1-400
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lane-batch 32 > gpurun_out/bench_stt_l32.log 2>&1; tail -1 gpurun_out/bench_stt_l32.log | cut -c1-400
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --batch 512 > gpurun_out/bench_stt_b512.log 2>&1; tail -1 gpurun_out/bench_stt_b512.log | cut -c1-400
