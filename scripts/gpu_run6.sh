mkdir -p gpurun_out
timeout 600 python scripts/profile_step.py > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"janus|gemm|attention|layernorm|logmel|dec_|yin|prosody" --csv --log-file gpurun_out/launches_step.csv python scripts/profile_step.py > gpurun_out/ncu_step.log 2>&1
tail -2 gpurun_out/ncu_step.log
