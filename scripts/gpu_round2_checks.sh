mkdir -p gpurun_out
timeout 900 python bench.py --model large-v3 --beam 5 --batch 256 --lane-batch 256 --steps 3 --warmup 3 --ref-windows 1 --parity-windows 0 > gpurun_out/r2_bench_beam5_largev3_b256.json 2> gpurun_out/r2_bench_beam5_largev3_b256.err; echo lv3 rc=$?; tail -c 700 gpurun_out/r2_bench_beam5_largev3_b256.json
timeout 300 python scripts/gpu_gemm_bench.py > gpurun_out/r2_gemm_bench.log 2>&1; cat gpurun_out/r2_gemm_bench.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 8 -c 8 -o gpurun_out/r2_gemm_shapes python scripts/gpu_gemm_bench.py > gpurun_out/r2_gemm_ncu.log 2>&1; echo ncu rc=$?
timeout 600 python -m pytest tests/test_gpu_transcribe.py tests/test_gpu_service.py tests/test_gpu_multidevice.py -q > gpurun_out/r2_gputests4.log 2>&1; echo tests rc=$?; tail -4 gpurun_out/r2_gputests4.log
timeout 300 python scripts/latency_transcriber.py > gpurun_out/r2_latency.json 2> gpurun_out/r2_latency.err; echo latency rc=$?; grep -A4 stt_ms gpurun_out/r2_latency.json | head -12
