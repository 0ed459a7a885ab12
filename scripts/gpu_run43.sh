mkdir -p gpurun_out
python -m janus_b200.build --force > /dev/null 2>&1
timeout 600 python __graft_entry__.py --smoke 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n2_latent.log 2>&1; tail -1 gpurun_out/bench_n2_latent.log | cut -c1-220
timeout 600 python bench_stream.py --sessions 1024 > gpurun_out/stream_n1_latent.log 2>&1; tail -1 gpurun_out/stream_n1_latent.log | cut -c1-900
