import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
_abi.lib()
d = 768
bar = torch.zeros(2, dtype=torch.int64, device="cuda")
pos = torch.zeros(1, dtype=torch.int32, device="cuda")
lw = torch.randn(d, device="cuda"); lb = torch.randn(d, device="cuda")
rows = 256
for (N, K, cfg, ln) in ((768, 768, 0, 0), (2304, 768, 1, 1), (3072, 768, 2, 0)):
    A = torch.randn(rows, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** .5).bfloat16()
    bias = torch.randn(N, device="cuda"); out = torch.empty(rows, N, device="cuda", dtype=torch.bfloat16)
    _abi.call("janus_debug_fused_linear", A.data_ptr(), W.data_ptr(), bias.data_ptr(), out.data_ptr(),
              lw.data_ptr() if ln else None, lb.data_ptr() if ln else None, rows, N, K, d, cfg, 8, bar.data_ptr(), pos.data_ptr(),
              torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
print("ok")
