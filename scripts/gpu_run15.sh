mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
B=64 timeout 300 python scripts/gpu_dec_bench.py 2>&1 | tail -12
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_stt_l64.log 2>&1; tail -1 gpurun_out/bench_stt_l64.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms'], d['roofline']['frac'])"
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lane-batch 32 > gpurun_out/bench_stt_l32.log 2>&1; tail -1 gpurun_out/bench_stt_l32.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms'], d['roofline']['frac'])"
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --batch 512 > gpurun_out/bench_stt_b512.log 2>&1; tail -1 gpurun_out/bench_stt_b512.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms'], d['roofline']['frac'])"
