mkdir -p gpurun_out
for c in 64 128 256; do
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --enc-chunk $c > gpurun_out/bench_c$c.log 2>&1; tail -1 gpurun_out/bench_c$c.log | python -c "
import json,sys
try:
    j=json.loads(sys.stdin.read()); print($c, round(j['value']), round(j['ms_per_step'],1), j['stages_ms']['logmel_encoder_ms'], j['clocks']['sm_mhz'])
except Exception as e: print('fail', e)"
done
