"""K1/K2 parity on a real B200 against the CPU oracles (oracle/logmel_ref.py, oracle/whisper_ref.py).

Stated tolerances
  log-mel (fp32 kernel vs numpy oracle): |diff| <= 5e-3 everywhere, mean |diff| <= 2e-5
      (values are (log10 P + 4)/4; fp32 FFT rounding dominates in bins ~80 dB below the frame peak).
  layernorm / attention building blocks (bf16 I/O): |err| <= 2e-2 * max(1, |ref|).
  encoder states (bf16 kernels vs fp32 oracle, output is layer-normed, unit scale):
      mean |err| <= 1.5e-2, max |err| <= 0.25.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _stream():
    return torch.cuda.current_stream().cuda_stream


def test_layernorm_block():
    from janus_b200 import _abi
    torch.manual_seed(0)
    for rows, d in ((1000, 768), (37, 128), (513, 1280), (64, 384)):
        x = (torch.randn(rows, d, device="cuda") * 3 + 1).bfloat16()
        w, b = torch.randn(d, device="cuda"), torch.randn(d, device="cuda")
        y = torch.empty_like(x)
        _abi.call("janus_layernorm_bf16", x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), rows, d, _stream())
        ref = torch.nn.functional.layer_norm(x.float(), (d,), w, b, 1e-5)
        err = (y.float() - ref).abs()
        assert bool((err <= 2e-2 * ref.abs().clamp(min=1)).all()), err.max()


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (2, 300, 2), (1, 1500, 6), (3, 1500, 12)])
def test_attention_block(B, T, H):
    from janus_b200 import _abi
    torch.manual_seed(B * 1000 + T + H)
    d = 64 * H
    qkv = (torch.randn(B, T, 3 * d, device="cuda") * 1.5).bfloat16()
    out = torch.zeros(B, T, d, device="cuda", dtype=torch.bfloat16)
    _abi.call("janus_attention_enc_bf16", qkv.data_ptr(), out.data_ptr(), B, T, H, _stream())
    torch.cuda.synchronize()
    q, k, v = (t.float().view(B, T, H, 64).permute(0, 2, 1, 3) for t in qkv.split(d, dim=-1))
    ref = torch.softmax(q @ k.transpose(-1, -2) / 8.0, dim=-1) @ v
    ref = ref.permute(0, 2, 1, 3).reshape(B, T, d)
    err = (out.float() - ref).abs()
    assert bool((err <= 2e-2 * ref.abs().clamp(min=1)).all()), f"max err {err.max().item()}"


def _clips():
    from janus_b200.synth import make_clips
    c = make_clips(4, 30.0, 16000, seed=77).numpy()
    return [c[0], c[1][:80000], c[2][:16000 * 11 + 37], c[3][:4000]]


@pytest.mark.parametrize("n_mels", [80, 128])
def test_logmel_vs_oracle(n_mels):
    from janus_b200.whisper.frontend import LogMelFrontend
    from oracle.logmel_ref import log_mel_window
    dev = torch.device("cuda", 0)
    clips = _clips()
    fe = LogMelFrontend(n_mels, dev)
    lens = torch.tensor([len(c) for c in clips], dtype=torch.int64)
    offs = torch.cumsum(lens, 0) - lens
    pcm = torch.from_numpy(np.concatenate(clips)).to(dev)
    mel_pad, mel = fe(pcm, offs.to(dev), lens.to(dev), decim=1, want_bf16=True, want_f32=True)
    mel = mel.cpu().numpy()
    for i, c in enumerate(clips):
        ref = log_mel_window(c, n_mels)
        diff = np.abs(mel[i] - ref)
        assert diff.max() <= 5e-3 and diff.mean() <= 2e-5, (i, diff.max(), diff.mean())
        n_content = (len(c) + 160) // 160 - 1
        assert (mel[i][:, min(n_content, 3000):] == 0).all()
    mp = mel_pad.float().cpu().numpy()
    assert (mp[:, 0] == 0).all() and (mp[:, 3001] == 0).all()
    assert np.abs(mp[:, 1:3001].transpose(0, 2, 1) - mel).max() <= 1e-2  # bf16 rounding of O(1) values


def test_logmel_decimation_fused():
    """48 kHz input with decim=3 == frontend of audio[::3] (transcriber.py:51)."""
    from janus_b200.synth import make_clips
    from janus_b200.whisper.frontend import LogMelFrontend
    dev = torch.device("cuda", 0)
    x48 = make_clips(1, 3.0, 48000, seed=5).to(dev)[0]
    fe = LogMelFrontend(80, dev)
    one = lambda n: torch.tensor([n], dtype=torch.int64, device=dev)  # noqa: E731
    _, a = fe(x48, one(0), one(x48.numel()), decim=3, want_bf16=False, want_f32=True)
    x16 = x48[::3].contiguous()
    _, b = fe(x16, one(0), one(x16.numel()), decim=1, want_bf16=False, want_f32=True)
    assert torch.equal(a, b)


@pytest.mark.parametrize("name", ["test", "tiny"])
def test_encoder_vs_fp32_oracle(name):
    from janus_b200.whisper.config import PRESETS, mini_dims
    from janus_b200.whisper.engine import WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = mini_dims(d=128, heads=2, layers=2) if name == "test" else PRESETS["tiny"]
    W = init_random_weights(dims, seed=3)
    clips = _clips()[:2]
    mel = torch.stack([torch.from_numpy(log_mel_window(c, dims.n_mels)) for c in clips])
    ref = R.encode(W, dims, mel)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    out = eng.encode(eng.mel_pad_from_f32(mel)).float().cpu()
    err = (out - ref).abs()
    print(f"encoder {dims.name}: mean err {err.mean():.4g} max err {err.max():.4g} ref std {ref.std():.3f}")
    assert err.mean() <= 1.5e-2 and err.max() <= 0.25


def test_large_v3_shape_end_to_end():
    """BASELINE configs[3] architecture family (d = 1280, 20 heads, 128 mel bins, 51866 tokens) at 1 layer:
    K1 log-mel with 128 filters, encoder vs the fp32 oracle, and greedy tokens from audio under confident
    weights. (The 32-layer model only repeats these blocks.)"""
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.engine import DecodeSession, WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = mini_dims(d=1280, heads=20, layers=1, n_mels=128, vocab=51866)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=4, profile="confident")
    clips = _clips()[:2]
    mel = torch.stack([torch.from_numpy(log_mel_window(c, dims.n_mels)) for c in clips])
    ref = R.encode(W, dims, mel)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    dev = eng.device
    lens = torch.tensor([len(c) for c in clips], dtype=torch.int64)
    offs = torch.cumsum(lens, 0) - lens
    pcm = torch.from_numpy(np.concatenate(clips)).to(dev)
    mel_pad, _ = eng.frontend(pcm, offs.to(dev), lens.to(dev))
    enc = eng.encode(mel_pad)
    err = (enc.float().cpu() - ref).abs()
    print(f"large-v3-shaped encoder: mean err {err.mean():.4g} max err {err.max():.4g}")
    assert err.mean() <= 1.5e-2 and err.max() <= 0.3
    n_new = 12
    ref_tok, _ = R.greedy_decode(W, dims, spec, ref, n_new)
    ses = DecodeSession(eng, batch=2, max_new_tokens=n_new)
    toks, _ = ses.decode(enc)
    torch.cuda.synchronize()
    assert torch.equal(toks.cpu().long(), ref_tok)
