mkdir -p gpurun_out
for b in 2048 4096; do
timeout 1200 python bench.py --batch $b --lane-batch $b --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b$b.log 2>&1; tail -1 gpurun_out/bench_b$b.log | python -c "
import json,sys
try:
    j=json.loads(sys.stdin.read()); print($b, j['value'], j['ms_per_step'], j['e2e']['value'], j['stages_ms'], j['roofline']['frac'], j['clocks'])
except Exception as e: print('fail', e)"
tail -3 gpurun_out/bench_b$b.log | cut -c1-300 | grep -i "error\|memory" | head -3
done
