for ns in 0 400 800 1200; do
  if [ $ns = 0 ]; then python -m janus_b200.build --force > /dev/null 2>&1; else JANUS_NVCC_EXTRA=-DATT_STAGGER_NS=$ns python -m janus_b200.build --force > /dev/null 2>&1; fi
  echo "stagger $ns: $(timeout 300 python scripts/gpu_att_profile.py 2>&1 | grep kernel)"
done
