mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python scripts/gpu_gemm_bench.py 2>&1 | tail -8
timeout 900 python bench.py --steps 3 --no-cpu-baseline > gpurun_out/bench_a.log 2>&1; tail -1 gpurun_out/bench_a.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(round(d['value']), round(d['ms_per_step'],1), round(d['e2e']['value']), {k: round(v,1) for k,v in d['stages_ms'].items()}, round(d['roofline']['frac'],3), d['roofline']['encoder_tensor']['frac'])" || tail -5 gpurun_out/bench_a.log
