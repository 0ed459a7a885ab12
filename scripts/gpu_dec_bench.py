"""Micro-benchmarks of the decode-step building blocks at Whisper-small shapes (B=64)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi

B = int(os.environ.get("B", "64"))
d = 768

def timeit(fn, n=20, graph=True):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    if graph:   # python/ctypes launch overhead (~5 us) would dominate: replay a captured graph instead
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        fn = g.replay
        fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us

def lin(N, K, L=24, f32=False, which="dec_linear"):
    A = torch.randn(B, K, device="cuda").bfloat16()
    Ws = [(torch.randn(N, K, device="cuda") / K ** .5).bfloat16() for _ in range(L)]   # distinct weights: no L2 reuse
    bias = torch.randn(N, device="cuda")
    out = torch.empty(B, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    def f():
        for W in Ws:
            if which == "dec_linear":
                _abi.call("janus_dec_linear_bf16", A.data_ptr(), K, W.data_ptr(), bias.data_ptr(), None, out.data_ptr(), B, N, K, 0, int(f32), torch.cuda.current_stream().cuda_stream)
            else:
                _abi.call("janus_gemm_bf16", A.data_ptr(), K, B * K, W.data_ptr(), K, out.data_ptr(), N, B * N, 0, B, 1, N, K, bias.data_ptr(), None, None, 0, torch.cuda.current_stream().cuda_stream)
    us = timeit(f, 10) / L
    print(f"{which} N={N} K={K}: {us:.2f} us/op  ({N*K*2/us/1e6:.2f} TB/s weight stream)")

if __name__ == "__main__":
    for which in ("dec_linear", "gemm"):
        lin(768, 768, which=which); lin(2304, 768, which=which); lin(3072, 768, which=which); lin(768, 3072, which=which)
    lin(51904, 768, L=4, f32=True)
    x = torch.randn(B, d, device="cuda").bfloat16(); y = torch.empty_like(x); w = torch.randn(d, device="cuda")
    def ln10():
        for _ in range(10):
            _abi.call("janus_layernorm_bf16", x.data_ptr(), w.data_ptr(), w.data_ptr(), y.data_ptr(), B, d, torch.cuda.current_stream().cuda_stream)
    print("layernorm", timeit(ln10, 50) / 10, "us")
    # full decode step through the engine (graph replay)
    from janus_b200.whisper.config import PRESETS
    from janus_b200.whisper.engine import WhisperEngine, DecodeSession
    from janus_b200.whisper.weights import init_random_weights
    dims = PRESETS["small"]
    eng = WhisperEngine(dims, init_random_weights(dims, seed=0, profile="confident"))
    ses = DecodeSession(eng, B, 64)
    enc = torch.randn(B, 1500, d, device="cuda").bfloat16()
    ses._ensure_graph(); ses.cross_kv(enc); ses.reset()
    for _ in range(10): ses.graph.replay()
    print("decode step (graph):", timeit(lambda: ses.graph.replay(), 40, graph=False), "us")
    ses.reset()
    print("decode step (eager):", timeit(lambda: ses._step(), 20, graph=False), "us")
