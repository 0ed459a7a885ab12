mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_default.log 2>&1; tail -1 gpurun_out/bench_default.log | cut -c1-1200
B=1024 timeout 300 python scripts/profile_decode_step.py > /dev/null 2>&1 && \
B=1024 timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:dec_cross_attn -c 1 -o gpurun_out/prof_xattn_b1024 python scripts/profile_decode_step.py > gpurun_out/ncu_xattn.log 2>&1; tail -1 gpurun_out/ncu_xattn.log
