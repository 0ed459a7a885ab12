mkdir -p gpurun_out
timeout 600 python bench_stream.py --sessions 1024 > gpurun_out/stream_n1.log 2>&1; tail -2 gpurun_out/stream_n1.log | cut -c1-900
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench_stream.py --sessions 1024 > gpurun_out/stream_n2.log 2>&1; tail -1 gpurun_out/stream_n2.log | cut -c1-900
timeout 600 python -m pytest tests/test_gpu_stt.py -x -q -k large_v3 2>&1 | tail -2
