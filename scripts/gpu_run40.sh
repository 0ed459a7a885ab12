mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r40.log 2>&1; tail -1 gpurun_out/bench_r40.log | cut -c1-300; tail -1 gpurun_out/bench_r40.log | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print(j['e2e']); print(j['stages_ms']); print({k:v for k,v in j['roofline'].items() if k not in ('note',)}); print(j['clocks'])"
