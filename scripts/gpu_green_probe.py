"""Probe: SM partitioning with CUDA green contexts. Runs the HBM-bound decode cross-attention on one
partition and the tensor-bound encoder GEMM on the other, alone and concurrently."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cuda.bindings import driver as cu
from janus_b200 import _abi

def chk(r):
    if isinstance(r, tuple):
        err, rest = r[0], r[1:]
    else:
        err, rest = r, ()
    if err != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(str(err))
    return rest[0] if len(rest) == 1 else rest

torch.cuda.init(); torch.zeros(1, device="cuda")
_abi.lib()
dev = chk(cu.cuDeviceGet(0))
res = chk(cu.cuDeviceGetDevResource(dev, cu.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
print("device SMs:", res.sm.smCount)
N_ENC = int(os.environ.get("ENC_SMS", "96"))
out = chk(cu.cuDevSmResourceSplitByCount(1, res, 0, N_ENC))
groups, nb, rem = out
print("split:", nb, "group SMs", groups[0].sm.smCount, "remaining", rem.sm.smCount)
def mk(resource):
    desc = chk(cu.cuDevResourceGenerateDesc([resource], 1))
    g = chk(cu.cuGreenCtxCreate(desc, dev, cu.CUgreenCtxCreate_flags.CU_GREEN_CTX_DEFAULT_STREAM))
    s = chk(cu.cuGreenCtxStreamCreate(g, cu.CUstream_flags.CU_STREAM_NON_BLOCKING, 0))
    return g, s
gA, sA = mk(groups[0]); gB, sB = mk(rem)
stA = torch.cuda.ExternalStream(int(sA)); stB = torch.cuda.ExternalStream(int(sB))

# workloads
B, H, T, kv_ld, d = 256, 12, 1500, 1536, 768
q = torch.randn(B, d, device="cuda").bfloat16(); att = torch.empty_like(q)
ck = torch.randn(B, H, 64, kv_ld, device="cuda").bfloat16(); cv = torch.randn(B, H, T, 64, device="cuda").bfloat16()
M, N, K = 96000, 2304, 768
A = torch.randn(M, K, device="cuda").bfloat16(); W = (torch.randn(N, K, device="cuda") / 28).bfloat16()
o = torch.empty(M, N, device="cuda", dtype=torch.bfloat16); bias = torch.randn(N, device="cuda")
def xattn(st, n=40):
    for _ in range(n):
        _abi.call("janus_dec_cross_attn_bf16", q.data_ptr(), ck.data_ptr(), cv.data_ptr(), att.data_ptr(), B, H, T, kv_ld, st.cuda_stream)
def gemm(st, n=20):
    for _ in range(n):
        _abi.call("janus_gemm_bf16", A.data_ptr(), K, M * K, W.data_ptr(), K, o.data_ptr(), N, M * N, 0, M, 1, N, K, bias.data_ptr(), None, None, 0, st.cuda_stream)
def timed(fn_list):
    torch.cuda.synchronize()
    evs = []
    for st, fn in fn_list:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); evs.append((e0, e1, st))
    for (st, fn) in fn_list: fn(st)
    for (e0, e1, st) in evs: e1.record(st)
    torch.cuda.synchronize()
    return [e0.elapsed_time(e1) for (e0, e1, st) in evs]
main = torch.cuda.current_stream()
for name, sx, sg in (("primary ctx (all SMs)", main, main), ("green: xattn on B, gemm on A", stB, stA)):
    xattn(sx, 3); gemm(sg, 3); torch.cuda.synchronize()
    tx = timed([(sx, xattn)])[0]; tg = timed([(sg, gemm)])[0]
    print(f"{name}: xattn alone {tx/40*1e3:.1f} us/launch ({B*H*2*T*64*2/(tx/40)/1e6:.0f} GB/s), gemm alone {tg/20*1e3:.1f} us ({2*M*N*K/(tg/20)/1e9:.0f} TFLOP/s)")
    if sx is not sg:
        both = timed([(sx, xattn), (sg, gemm)])
        print(f"   concurrent: xattn {both[0]/40*1e3:.1f} us/launch, gemm {both[1]/20*1e3:.1f} us/launch")
s2 = torch.cuda.Stream()
both = timed([(main, xattn), (s2, gemm)])
print(f"two plain streams concurrent: xattn {both[0]/40*1e3:.1f} us/launch, gemm {both[1]/20*1e3:.1f} us/launch")
