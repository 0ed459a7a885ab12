mkdir -p gpurun_out
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.log 2>&1; tail -1 gpurun_out/bench_n$N.log | python -c "
import json,sys
try:
    j=json.loads(sys.stdin.read()); print('n_gpus', j['n_gpus'], 'value', round(j['value']), 'ms/step', round(j['ms_per_step'],1), 'e2e', round(j['e2e']['value']), j['clocks'])
except Exception as e: print('fail', e)"
tail -5 gpurun_out/bench_n$N.log | grep -i "error\|Traceback" | head -3
