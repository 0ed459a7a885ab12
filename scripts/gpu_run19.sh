mkdir -p gpurun_out
show() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('$1', round(d['value']), round(d['ms_per_step'],1), round(d['e2e']['value']), {k: round(v,1) for k,v in d['stages_ms'].items()}, round(d['roofline']['frac'],3))" || tail -5 $1; }
timeout 300 python -m pytest tests/test_gpu_decoder.py tests/test_gpu_service.py -x -q 2>&1 | tail -3
for cfg in "1024 1024 2" "1024 512 4" "512 512 2"; do
set -- $cfg
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --batch $1 --lane-batch $2 --lanes $3 > gpurun_out/bench_b$1_l$2x$3.log 2>&1; show gpurun_out/bench_b$1_l$2x$3.log
nvidia-smi --query-gpu=memory.used --format=csv,noheader
done
