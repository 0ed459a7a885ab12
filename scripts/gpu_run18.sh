mkdir -p gpurun_out
show() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('$1', round(d['value']), round(d['ms_per_step'],1), round(d['e2e']['value']), {k: round(v,1) for k,v in d['stages_ms'].items()}, round(d['roofline']['frac'],3))" || tail -5 $1; }
for cfg in "256 2" "128 4" "64 8"; do
set -- $cfg
timeout 900 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --lane-batch $1 --lanes $2 > gpurun_out/bench_f$1x$2.log 2>&1; show gpurun_out/bench_f$1x$2.log
done
