"""One launch each of the encoder's four GEMM shapes, the encoder attention and the latent cross-attention at the
headline shapes, inside a cudaProfilerStart/Stop range, for `ncu --set full --profile-from-start off`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi

st = torch.cuda.current_stream().cuda_stream
M = 64 * 1500


def gemm(N, K, act=0, res=False):
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    b = torch.randn(N, device="cuda")
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    r = torch.randn(M, N, device="cuda").bfloat16() if res else None
    return lambda: _abi.call("janus_gemm_bf16", A.data_ptr(), K, M * K, W.data_ptr(), K, out.data_ptr(), N, M * N, 0, M, 1, N,
                             K, b.data_ptr(), r.data_ptr() if res else None, None, act, st), (A, W, b, out, r)


jobs = [gemm(768, 768, res=True), gemm(2304, 768), gemm(3072, 768, act=1), gemm(768, 3072, res=True)]
B, T, H = 64, 1500, 12
qkv = torch.randn(B, T, 3 * H * 64, device="cuda").bfloat16()
o = torch.empty(B, T, H * 64, device="cuda", dtype=torch.bfloat16)
jobs.append((lambda: _abi.call("janus_attention_enc_bf16", qkv.data_ptr(), o.data_ptr(), B, T, H, st), None))
n, d = 1024, 768
enc = torch.randn(n, T, d, device="cuda").bfloat16()
qlat = torch.zeros(n, 16, d, device="cuda", dtype=torch.bfloat16)
qlat[:, :H] = (torch.randn(n, H, d, device="cuda") * (8.0 / d ** 0.5)).bfloat16()
z = torch.empty(n, H, d, device="cuda", dtype=torch.bfloat16)
jobs.append((lambda: _abi.call("janus_dec_xlat_bf16", qlat.data_ptr(), enc.data_ptr(), z.data_ptr(), n, T, d, H, st), None))
for f, _ in jobs:
    for _ in range(3):
        f()
torch.cuda.synchronize()
torch.cuda.profiler.start()
for f, _ in jobs:
    f()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
