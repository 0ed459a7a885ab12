import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
_abi.lib()
d = 768
bar = torch.zeros(2, dtype=torch.int64, device="cuda")
pos = torch.zeros(1, dtype=torch.int32, device="cuda")
lw = torch.randn(d, device="cuda"); lb = torch.randn(d, device="cuda")
def t(rows, N, K, cfg, ln, nph, reps=20):
    A = torch.randn(rows, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** .5).bfloat16()
    bias = torch.randn(N, device="cuda"); out = torch.empty(rows, N, device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    def f():
        _abi.call("janus_debug_fused_linear", A.data_ptr(), W.data_ptr(), bias.data_ptr(), out.data_ptr(),
                  lw.data_ptr() if ln else None, lb.data_ptr() if ln else None, rows, N, K, d, cfg, nph, bar.data_ptr(), pos.data_ptr(), st)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
for rows in (64, 256):
    for (name, N, K, cfg, ln) in (("d->d cfg0", 768, 768, 0, 0), ("LN d->d cfg0", 768, 768, 0, 1), ("LN qkv cfg1", 2304, 768, 1, 1),
                                  ("LN fc1 cfg2", 3072, 768, 2, 1), ("fc2 cfg0", 768, 3072, 0, 0)):
        t1, t8 = t(rows, N, K, cfg, ln, 1), t(rows, N, K, cfg, ln, 8)
        print(f"rows={rows} {name}: 1 phase {t1:.1f} us, 8 phases {t8:.1f} us -> {(t8 - t1) / 7:.2f} us per phase+barrier")
