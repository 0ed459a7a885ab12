mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_latent.log 2>&1; tail -1 gpurun_out/bench_latent.log | cut -c1-200
B=1024 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_dec1024_latent.csv python scripts/profile_decode_step.py > gpurun_out/ncu_l.log 2>&1; tail -1 gpurun_out/ncu_l.log
B=1024 timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:dec_xlat -c 1 -f -o gpurun_out/prof_xlat_b1024 python scripts/profile_decode_step.py > gpurun_out/ncu_xlat.log 2>&1; tail -1 gpurun_out/ncu_xlat.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_enc -s 2 -c 1 -f -o gpurun_out/prof_attn5 python scripts/gpu_att_once.py > gpurun_out/ncu_attn5.log 2>&1; tail -1 gpurun_out/ncu_attn5.log
