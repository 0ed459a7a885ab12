mkdir -p gpurun_out
export B=64 NEW=4 REPS=2 LANE=64
timeout 600 python scripts/profile_step.py > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attention_enc" -s 14 -c 1 -o gpurun_out/prof_attn2 python scripts/profile_step.py > gpurun_out/ncu_full_attn.log 2>&1
tail -2 gpurun_out/ncu_full_attn.log
