JANUS_NVCC_EXTRA="-DATT_PROFILE -DATT_POLY_MODE=3" python -m janus_b200.build --force 2>&1 | tail -1
timeout 300 python scripts/gpu_att_profile.py 2>&1 | tail -12
