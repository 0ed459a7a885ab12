"""Decode-chain micro-benchmarks: PDL on/off for chains of skinny linears / layernorms / full step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi

B = int(os.environ.get("B", "64"))
d = 768
L_ = _abi.lib()

def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us

def lin_chain(N, K, L=24, f32=False):
    A = torch.randn(B, K, device="cuda").bfloat16()
    Ws = [(torch.randn(N, K, device="cuda") / K ** .5).bfloat16() for _ in range(L)]
    bias = torch.randn(N, device="cuda")
    out = torch.empty(B, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    def f():
        st = torch.cuda.current_stream().cuda_stream
        for W in Ws:
            _abi.call("janus_dec_linear_bf16", A.data_ptr(), K, W.data_ptr(), bias.data_ptr(), None, out.data_ptr(), B, N, K, 0, int(f32), st)
    return timeit(f, 10) / L

def ln_chain(rows, L=40):
    x = torch.randn(rows, d, device="cuda").bfloat16(); y = torch.empty_like(x); w = torch.randn(d, device="cuda")
    def f():
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(L):
            _abi.call("janus_layernorm_bf16", x.data_ptr(), w.data_ptr(), w.data_ptr(), y.data_ptr(), rows, d, st)
    return timeit(f, 20) / L

if __name__ == "__main__":
    for pdl in (0, 1):
        L_.janus_set_pdl(pdl)
        print(f"--- pdl={pdl}  B={B}")
        for (N, K) in ((768, 768), (2304, 768), (3072, 768), (768, 3072)):
            print(f"dec_linear N={N} K={K}: {lin_chain(N, K):.2f} us/op")
        print(f"dec_linear N=51904 K=768: {lin_chain(51904, 768, L=4, f32=True):.2f} us/op")
        print(f"layernorm rows={B}: {ln_chain(B):.2f} us   rows=8: {ln_chain(8):.2f} us")
    from janus_b200.whisper.config import PRESETS
    from janus_b200.whisper.engine import WhisperEngine, DecodeSession
    from janus_b200.whisper.weights import init_random_weights
    dims = PRESETS["small"]
    eng = WhisperEngine(dims, init_random_weights(dims, seed=0, profile="confident"))
    enc = torch.randn(B, 1500, d, device="cuda").bfloat16()
    for pdl in (0, 1):
        L_.janus_set_pdl(pdl)
        ses = DecodeSession(eng, B, 64, light=True, steps_per_graph=10)
        ses._ensure_graph(); ses.cross_kv(enc); ses.reset()
        ses.graph.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(4): ses.graph.replay()
        e1.record(); torch.cuda.synchronize()
        print(f"pdl={pdl}: light decode step (10-step graph): {e0.elapsed_time(e1) / 40 * 1e3:.1f} us")
