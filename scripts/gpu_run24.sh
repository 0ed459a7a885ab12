mkdir -p gpurun_out
export B=64 NEW=1 REPS=1 LANE=64
timeout 600 python scripts/profile_step.py > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm|attention|layernorm|logmel|yin|prosody" --csv --log-file gpurun_out/launches_enc64.csv python scripts/profile_step.py > gpurun_out/ncu_enc.log 2>&1
tail -1 gpurun_out/ncu_enc.log
