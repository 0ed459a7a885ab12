"""Latent cross-attention kernel (janus_b200/csrc/dec_xlat.cu) against a float32 torch restatement of the same
contraction: z[w, h] = softmax_t(X[w] . q'[w, h] / 8) @ X[w]. The kernel rounds p to bf16 for the second GEMM
and accumulates in fp32, so the tolerance is bf16-sized (2e-2 of the row scale)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(n, T, d, H, seed, sharp=1.0):
    from janus_b200 import _abi
    g = torch.Generator(device="cuda").manual_seed(seed)
    enc = torch.randn(n, T, d, device="cuda", generator=g).bfloat16()
    qlat = torch.zeros(n, 16, d, device="cuda", dtype=torch.bfloat16)
    qlat[:, :H] = (torch.randn(n, H, d, device="cuda", generator=g) * (sharp * 8.0 / d ** 0.5)).bfloat16()
    z = torch.full((n, H, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    _abi.call("janus_dec_xlat_bf16", qlat.data_ptr(), enc.data_ptr(), z.data_ptr(), n, T, d, H, st)
    torch.cuda.synchronize()
    X = enc.float()
    s = torch.einsum("ntd,nhd->nht", X, qlat[:, :H].float()) * 0.125
    ref = torch.einsum("nht,ntd->nhd", torch.softmax(s, dim=-1), X)
    return z.float(), ref


@pytest.mark.parametrize("n,T,d,H", [(3, 1500, 768, 12), (2, 100, 384, 6), (5, 200, 384, 6), (4, 256, 512, 8),
                                     (310, 300, 768, 12), (150, 700, 1024, 16)])
def test_xlat_matches_float32(n, T, d, H):
    z, ref = _run(n, T, d, H, seed=7 + n)
    assert torch.isfinite(z).all()
    scale = ref.abs().amax(dim=-1, keepdim=True).clamp_min(1e-3)
    err = ((z - ref).abs() / scale).max().item()
    assert err < 2e-2, err


def test_xlat_streaming_variant_both_in_one_process():
    """d = 1024 (Whisper-medium) does not fit the one-tile ring and runs the streaming two-pass kernel; the same
    kernel is selected for small d with janus_set_option(JANUS_OPT_XLAT_STREAM, 1) -- an explicit call, so both variants
    are exercised here in one process and must agree with float32 (and closely with each other)."""
    from janus_b200 import _abi
    z, ref = _run(3, 1500, 1024, 16, seed=5)
    scale = ref.abs().amax(dim=-1, keepdim=True).clamp_min(1e-3)
    assert ((z - ref).abs() / scale).max().item() < 2e-2
    za, ref = _run(5, 1500, 768, 12, seed=9)
    try:
        _abi.set_option(_abi.OPT_XLAT_STREAM, 1)
        zb, _ = _run(5, 1500, 768, 12, seed=9)
    finally:
        _abi.set_option(_abi.OPT_XLAT_STREAM, -1)
    scale = ref.abs().amax(dim=-1, keepdim=True).clamp_min(1e-3)
    assert ((za - ref).abs() / scale).max().item() < 2e-2 and ((zb - ref).abs() / scale).max().item() < 2e-2
    assert ((za - zb).abs() / scale).max().item() < 2e-2
    zc, _ = _run(5, 1500, 768, 12, seed=9)
    assert torch.equal(za, zc)          # back on the default kernel: bit-identical to the first run


def test_xlat_sharp_distribution_rescales():
    # large score spread: the running maximum moves by more than the lazy-rescale threshold between tiles
    z, ref = _run(4, 1500, 768, 12, seed=3, sharp=12.0)
    assert torch.isfinite(z).all()
    scale = ref.abs().amax(dim=-1, keepdim=True).clamp_min(1e-3)
    assert ((z - ref).abs() / scale).max().item() < 3e-2
