# ncu evidence for the headline workload: launch list of one un-graphed pipeline step + full captures of the top kernels
mkdir -p gpurun_out
export B=64 NEW=4 REPS=2
timeout 600 python scripts/profile_step.py > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm|attention|layernorm|logmel|dec_|yin|prosody" --csv --log-file gpurun_out/launches_step.csv python scripts/profile_step.py > gpurun_out/ncu_step.log 2>&1
timeout 600 python scripts/profile_step.py > gpurun_out/plain_step2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dec_cross_attn|attention_enc" -s 2 -c 2 -o gpurun_out/prof_attn python scripts/profile_step.py > gpurun_out/ncu_full1.log 2>&1
timeout 600 python scripts/profile_step.py > gpurun_out/plain_step3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16_kernel" -s 4 -c 2 -o gpurun_out/prof_gemm python scripts/profile_step.py > gpurun_out/ncu_full2.log 2>&1
ls -la gpurun_out | tail -8
