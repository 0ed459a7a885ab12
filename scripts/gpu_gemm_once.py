"""One launch of each encoder-layer GEMM shape (for ncu --set full -k regex:gemm_bf16)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
M = 64 * 1500
st = torch.cuda.current_stream().cuda_stream
def run(N, K, act, res):
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    b = torch.randn(N, device="cuda")
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    r = torch.randn(M, N, device="cuda").bfloat16() if res else None
    for _ in range(2):
        _abi.call("janus_gemm_bf16", A.data_ptr(), K, M * K, W.data_ptr(), K, out.data_ptr(), N, M * N, 0, M, 1, N, K,
                  b.data_ptr(), r.data_ptr() if res else None, None, act, st)
    torch.cuda.synchronize()
run(768, 768, 0, True)
run(2304, 768, 0, False)
run(3072, 768, 1, False)
run(768, 3072, 0, True)
