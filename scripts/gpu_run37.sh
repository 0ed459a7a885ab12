mkdir -p gpurun_out
for m in 3 2 0 3; do
  JANUS_NVCC_EXTRA=-DATT_POLY_MODE=$m python -m janus_b200.build --force > /dev/null 2>&1
  echo "poly mode $m: $(timeout 300 python scripts/gpu_att_profile.py 2>&1 | grep kernel)"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_enc -s 2 -c 1 -f -o gpurun_out/prof_attn4_poly3 python scripts/gpu_att_once.py > gpurun_out/ncu_attn4.log 2>&1; tail -1 gpurun_out/ncu_attn4.log
JANUS_NVCC_EXTRA=-DATT_POLY_MODE=0 python -m janus_b200.build --force > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_enc -s 2 -c 1 -f -o gpurun_out/prof_attn4_poly0 python scripts/gpu_att_once.py > gpurun_out/ncu_attn4b.log 2>&1; tail -1 gpurun_out/ncu_attn4b.log
