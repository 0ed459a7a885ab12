"""Times dec_cross_attn_beam_kernel (beam-shared K/V cross-attention) alone: algorithmic GB/s per launch.
JANUS_XB_VARIANT selects the loads-in-flight variant (read once per process): run one process per variant.

    python scripts/gpu_xb_bench.py [n_streams] [n_heads] [beam]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from janus_b200 import _abi  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    H = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    beam = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    T, kv_ld, d = 1500, 1536, 64 * H
    rows = n * beam
    q = torch.randn(rows, d, device="cuda").bfloat16()
    L = 4   # distinct layers so that consecutive launches do not hit L2
    layout = int(os.environ.get("KV_LAYOUT", "1"))   # 1: K row-major / V transposed (tensor-core kernel)
    if layout:
        kT = torch.randn(L, n, H, T, 64, device="cuda").bfloat16()
        v = torch.randn(L, n, H, 64, kv_ld, device="cuda").bfloat16()
    else:
        kT = torch.randn(L, n, H, 64, kv_ld, device="cuda").bfloat16()
        v = torch.randn(L, n, H, T, 64, device="cuda").bfloat16()
    out = torch.empty(rows, d, device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream()

    def run(l):
        _abi.call("janus_dec_cross_attn_beam_bf16", q.data_ptr(), kT[l].data_ptr(), v[l].data_ptr(), out.data_ptr(), n,
                  beam, H, T, kv_ld, layout, st.cuda_stream)

    for l in range(L):
        run(l)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    a.record()
    for _ in range(reps):
        for l in range(L):
            run(l)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / (reps * L)
    alg = n * H * 2 * T * 64 * 2 + 2 * rows * d * 2
    print(f"kv_layout={layout} n_streams={n} H={H} beam={beam}: {ms:.4f} ms per launch, "
          f"{alg / ms / 1e6:.0f} GB/s algorithmic")


if __name__ == "__main__":
    main()
