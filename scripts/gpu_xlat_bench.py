"""Times the latent cross-attention kernel at the bench shape (1024 windows x 1500 states x 768)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
n, T, d, H = int(os.environ.get("N", 1024)), 1500, 768, 12
enc = torch.randn(n, T, d, device="cuda").bfloat16()
qlat = torch.zeros(n, 16, d, device="cuda", dtype=torch.bfloat16)
qlat[:, :H] = (torch.randn(n, H, d, device="cuda") * (8.0 / d ** 0.5)).bfloat16()
z = torch.empty(n, H, d, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.current_stream().cuda_stream
f = lambda: _abi.call("janus_dec_xlat_bf16", qlat.data_ptr(), enc.data_ptr(), z.data_ptr(), n, T, d, H, st)
for _ in range(3): f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): f()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
byt = enc.numel() * 2 + qlat.numel() * 2 + z.numel() * 2
print(f"dec_xlat n={n}: {ms:.3f} ms  {byt / ms / 1e6:.0f} GB/s  ({4.0 * n * T * d * 16 / ms / 1e9:.0f} TFLOP/s issued)")
try:   # phase timing, only in a -DXL_PROFILE build
    import ctypes
    fn = _abi.lib().janus_debug_xlat_profile
    buf = (ctypes.c_longlong * 16)()
    fn.restype = ctypes.c_int32
    assert fn(buf) == 0
    g = max(1, buf[5])
    print(f"MMA warp per tile: wait box0 {buf[0]/g:.0f}  wait other boxes {buf[1]/g:.0f}  gemm1 issue {buf[2]/g:.0f}  wait P {buf[3]/g:.0f}  gemm2 issue {buf[4]/g:.0f}  (tiles {g})")
    g = max(1, buf[12])
    print(f"softmax per tile: wait S {buf[8]/g:.0f}  ld+max+barrier {buf[9]/g:.0f}  rescale+p+publish {buf[10]/g:.0f}  epilogue/tile {buf[11]/g:.0f}")
except AttributeError:
    pass
