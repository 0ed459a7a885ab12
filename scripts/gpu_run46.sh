for sp in 1 2; do
echo "== split $sp"
JANUS_ATT_SPLIT=$sp timeout 600 python -m pytest tests/test_gpu_stt.py -x -q 2>&1 | tail -2
JANUS_ATT_SPLIT=$sp timeout 300 python scripts/gpu_att_profile.py 2>&1 | grep kernel
JANUS_ATT_SPLIT=$sp timeout 300 python scripts/gpu_gemm_bench.py 2>&1 | grep attention
done
