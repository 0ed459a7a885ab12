mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stt.py -x -q 2>&1 | tail -3
timeout 300 python scripts/gpu_gemm_bench.py 2>&1 | grep -i att
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_enc -s 2 -c 1 -f -o gpurun_out/prof_attn3 python scripts/gpu_att_once.py > gpurun_out/ncu_attn3.log 2>&1; tail -2 gpurun_out/ncu_attn3.log
JANUS_NVCC_EXTRA=-DATT_PROFILE python -m janus_b200.build --force 2>&1 | tail -1
timeout 300 python scripts/gpu_att_profile.py 2>&1 | tail -9
