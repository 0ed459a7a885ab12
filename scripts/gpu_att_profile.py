"""Per-phase cycle counts of attention_enc_kernel (needs a build with JANUS_NVCC_EXTRA=-DATT_PROFILE)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
L = _abi.lib()
B, T, H = 64, 1500, 12
torch.manual_seed(0)
qkv = torch.randn(B, T, 3 * H * 64, device="cuda").bfloat16()
o = torch.empty(B, T, H * 64, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    _abi.call("janus_attention_enc_bf16", qkv.data_ptr(), o.data_ptr(), B, T, H, st)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 72)()
try:
    fn = L.janus_debug_att_profile
    fn.restype = ctypes.c_int32
    assert fn(buf) == 0
except AttributeError:
    pass   # not an ATT_PROFILE build: only the kernel time below is meaningful
names = ["wait_S", "tmem_ld", "max+rescale", "exp", "wait_O+st", "fence+arrive", "total", "tiles"]
for w in range(8):
    row = [buf[w * 8 + i] for i in range(8)]
    if row[7] == 0: continue
    n = row[7]
    print(f"warp {w}: " + "  ".join(f"{nm}={row[i] / n:.0f}" for i, nm in enumerate(names[:6])) + f"  total/tile={row[6] / n:.0f}")
print(f"CTA: prologue={buf[64]} clk  loop+epilogue={buf[65]} clk  entry->exit={buf[66]} clk")
print(f"issuer: in issue_s={buf[67]} clk  issue_pv={buf[68]}  load_kv={buf[69]}  loop iterations={buf[70]}  loop total={buf[71]} clk")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    _abi.call("janus_attention_enc_bf16", qkv.data_ptr(), o.data_ptr(), B, T, H, st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
ctas = B * H * ((T + 255) // 256)
print(f"kernel {ms:.3f} ms; {ctas} CTAs -> {ms * 1e-3 / (ctas / 148) * 1e6:.2f} us per CTA slot")
