"""Times the tcgen05 GEMM at the Whisper-small encoder shapes next to torch.matmul (cuBLAS)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi

def run(M, N, K, act=0, bias=True, res=False, iters=20):
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    b = torch.randn(N, device="cuda") if bias else None
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    r = torch.randn(M, N, device="cuda").bfloat16() if res else None
    st = torch.cuda.current_stream().cuda_stream
    def f():
        _abi.call("janus_gemm_bf16", A.data_ptr(), K, M * K, W.data_ptr(), K, out.data_ptr(), N, M * N, 0, M, 1, N, K,
                  b.data_ptr() if bias else None, r.data_ptr() if res else None, None, act, st)
    def g():
        torch.matmul(A, W.T)
    res_ms = []
    for fn in (f, g):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        res_ms.append(e0.elapsed_time(e1) / iters)
    fl = 2.0 * M * N * K
    print(f"M={M} N={N} K={K} act={act} res={res}: janus {res_ms[0]:.3f} ms {fl/res_ms[0]/1e9:.0f} TFLOP/s | cuBLAS {res_ms[1]:.3f} ms {fl/res_ms[1]/1e9:.0f} TFLOP/s")

if __name__ == "__main__":
    M = 64 * 1500
    run(M, 768, 768, res=True)
    run(M, 2304, 768)
    run(M, 3072, 768, act=1)
    run(M, 768, 3072, res=True)
    run(M, 18432, 768, bias=False, iters=5)
    run(8192, 8192, 8192, bias=False, iters=5)
    # encoder attention at the micro-batch shape (64 windows x 1500 x 12 heads)
    B, T, H = 64, 1500, 12
    qkv = torch.randn(B, T, 3 * H * 64, device="cuda").bfloat16()
    o = torch.empty(B, T, H * 64, device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    fa = lambda: _abi.call("janus_attention_enc_bf16", qkv.data_ptr(), o.data_ptr(), B, T, H, st)
    for _ in range(3): fa()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fa()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"attention_enc B={B}: {ms:.3f} ms  {4.0 * T * T * 64 * H * B / ms / 1e9:.0f} TFLOP/s")
