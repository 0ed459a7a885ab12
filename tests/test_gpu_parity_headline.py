"""Oracle parity AT THE HEADLINE MODEL (Whisper-small, 12 + 12 layers, BASELINE configs[2]) and at full-depth tiny,
on the decode path the bench times: latent cross-attention (dec_xlat.cu) and, with a 128-window lane, the grouped
tcgen05 GEMM route of the batched engine.

What is asserted, against the fp32 CPU oracle (oracle/whisper_ref.py, oracle/logmel_ref.py) on the same audio and
the same (bf16-representable) random-init weights, "confident" profile:

  encoder states      from audio (K1 + K2): mean |err| <= 1.5e-2, max |err| <= 0.3 on the layer-normed output
  teacher-forced      the CUDA decoder is fed the ORACLE's 128 tokens; at every position whose fp32 top-1 / top-2 margin
                      exceeds MARGIN_TOL its rule-filtered argmax equals the oracle's token (all windows, all positions)
  free-running        128 greedy tokens from audio: a window whose minimum fp32 margin exceeds MARGIN_TOL is
                      BIT-EXACT; any other window is bit-exact up to its first divergence, and the oracle's margin at
                      that position is <= MARGIN_TOL (a near-tie that bf16 rounding may legitimately resolve either
                      way: north_star's "bit-exact greedy tokens" is only well-defined away from ties)

"Margin" is the oracle's DECISION margin of a greedy step (oracle/whisper_ref.greedy_decode): the smaller of the top-1 /
top-2 gap of the rule-filtered logits and the gap of the "timestamps outweigh text" rule decision itself
(|log sum p(timestamps) - max log p(text)|) -- a near-tie in that rule flips the token between the two categories even
when the gap inside either category is large (found by checking 32 windows of the bench batch: two windows diverged at
top-1 / top-2 gaps of 0.7 and 1.5 because the rule's own decision was a near-tie). MARGIN_TOL is stated in logit units
next to the measured bf16 logit error of the same run (printed). Every test prints the minimum oracle margin it ran at,
so the statement can be seen not to be vacuous.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

# measured on B200 (round 2): last-step logit |err| mean 0.008 / max 0.10 (small), 0.007 / 0.075 (tiny), logit std 1.1
MARGIN_TOL = {"small": 0.25, "tiny": 0.2}
N_NEW = 128


def _clips(k):
    from janus_b200.synth import make_clips
    c = make_clips(k, 30.0, 16000, seed=1234).numpy()   # the bench's windows 0 .. k-1 (same per-clip seeds)
    out = [c[i] for i in range(k)]
    if k > 1:
        out[1] = out[1][:16000 * 11 + 37]               # one ragged window (zero-padded in feature space)
    return out


_ORACLE = {}


def _oracle(name, k):
    """Oracle pipeline for the first k bench windows of preset `name` (cached per test session)."""
    key = (name, k)
    if key in _ORACLE:
        return _ORACLE[key]
    from janus_b200.whisper.config import PRESETS, tokenizer_spec
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = PRESETS[name]
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=0, profile="confident")   # SemanticEncoder / bench defaults
    clips = _clips(k)
    with torch.no_grad():
        mel = torch.stack([torch.from_numpy(log_mel_window(c, dims.n_mels)) for c in clips])
        enc = R.encode(W, dims, mel)
        tok, lp, margins = R.greedy_decode(W, dims, spec, enc, N_NEW, return_margins=True)
    _ORACLE[key] = (dims, spec, W, clips, enc, tok, lp, margins)
    return _ORACLE[key]


def check_free_running(got: torch.Tensor, ref: torch.Tensor, margins: torch.Tensor, tol: float, what: str):
    """The free-running statement of the module docstring. Returns (#bit-exact windows, margin at the first
    divergences)."""
    got, ref = got.cpu().long(), ref.cpu().long()
    exact, div_margins = 0, []
    for b in range(ref.shape[0]):
        m = margins[b]
        min_m = float(m[torch.isfinite(m)].min())
        neq = (got[b] != ref[b]).nonzero()
        if neq.numel() == 0:
            exact += 1
            print(f"{what} window {b}: bit-exact over {ref.shape[1]} tokens, min oracle margin {min_m:.3f}")
            continue
        t = int(neq[0])
        div_margins.append(float(m[t]))
        print(f"{what} window {b}: first divergence at token {t} where the oracle margin is {float(m[t]):.4f} "
              f"(window min margin {min_m:.3f})")
        assert min_m <= tol, f"{what} window {b} differs although its min margin {min_m} > {tol}"
        assert float(m[t]) <= tol, f"{what} window {b} diverges at token {t} at margin {float(m[t])} > {tol}"
    return exact, div_margins


def check_teacher_forced(choice: torch.Tensor, ref: torch.Tensor, margins: torch.Tensor, tol: float, what: str):
    choice, ref = choice.cpu().long(), ref.cpu().long()
    safe = margins > tol
    bad = (choice != ref) & safe
    agree = float((choice == ref).float().mean())
    fin = margins[torch.isfinite(margins)]
    print(f"{what}: teacher-forced argmax agreement {agree:.4f} over {ref.numel()} positions, "
          f"{int(safe.sum())} above margin {tol}; oracle margin min {float(fin.min()):.4f} median {float(fin.median()):.3f}; "
          f"largest margin at a disagreement "
          f"{float(margins[choice != ref].max()) if bool((choice != ref).any()) else 0.0:.4f}")
    assert not bool(bad.any()), f"{what}: argmax differs at margins {margins[bad].tolist()}"
    return agree


def _gpu_encode(eng, clips):
    dev = eng.device
    lens = torch.tensor([len(c) for c in clips], dtype=torch.int64)
    offs = torch.cumsum(lens, 0) - lens
    pcm = torch.from_numpy(np.concatenate(clips)).to(dev)
    mel_pad, _ = eng.frontend(pcm, offs.to(dev), lens.to(dev))
    return eng.encode(mel_pad)


@pytest.mark.parametrize("name,k", [("small", 3), ("tiny", 4)])
def test_full_depth_encoder_and_latent_greedy_vs_oracle(name, k):
    from janus_b200.whisper.engine import DecodeSession, WhisperEngine
    dims, spec, W, clips, enc_ref, ref_tok, ref_lp, margins = _oracle(name, k)
    tol = MARGIN_TOL[name]
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    enc = _gpu_encode(eng, clips)
    err = (enc.float().cpu() - enc_ref).abs()
    print(f"{name} encoder ({dims.n_audio_layer} layers) vs fp32 oracle: mean |err| {err.mean():.4g} max {err.max():.4g} "
          f"(ref std {enc_ref.std():.3f})")
    assert err.mean() <= 1.5e-2 and err.max() <= 0.3

    ses = DecodeSession(eng, batch=k, max_new_tokens=N_NEW, latent=True)
    assert ses.latent
    # (1) teacher-forced on the CUDA encoder states: every position of every window
    ses.decode(enc, forced=ref_tok.cuda())
    torch.cuda.synchronize()
    check_teacher_forced(ses.choice[:, ses.n_prompt:], ref_tok, margins, tol, f"{name} latent")
    # final-step logits (all 51865 columns) for scale: the tolerance above is in these units
    from oracle import whisper_ref as R
    st = R.DecoderState(W, dims, enc_ref)
    n_tok = [int((ref_tok[b] != spec.eot).sum()) for b in range(k)]
    if min(n_tok) == N_NEW:   # no window stopped early: the last step's logits exist for every row
        full = torch.cat([torch.tensor([list(spec.sot_sequence)] * k), ref_tok[:, :-1]], 1)
        with torch.no_grad():
            ref_logits = st.step(full)[:, -1]
        lerr = (ses.logits[:, :dims.n_vocab].cpu() - ref_logits).abs()
        print(f"{name} latent: last-step logit |err| mean {lerr.mean():.4f} max {lerr.max():.4f} (logit std {ref_logits.std():.2f})")
        assert lerr.mean() <= 0.06
    # (2) free-running from audio
    toks, sum_lp = ses.decode(enc)
    torch.cuda.synchronize()
    exact, _ = check_free_running(toks, ref_tok, margins, tol, f"{name} latent")
    print(f"{name} latent: {exact}/{k} windows bit-exact over {N_NEW} free-running tokens")
    # (3) the K/V-cache session satisfies the same statement
    kv = DecodeSession(eng, batch=k, max_new_tokens=N_NEW, latent=False)
    tk, _ = kv.decode(enc)
    torch.cuda.synchronize()
    check_free_running(tk, ref_tok, margins, tol, f"{name} K/V")


def test_small_engine_lane128_latent_grouped_gemm_vs_oracle():
    """The bench's own route: SemanticEncoder('small') with a 128-window lane (latent cross-attention chosen by
    the engine itself, decode GEMMs through the tcgen05 kernel at M = 128, grouped absorb / value GEMMs). Windows
    0 .. 2 are the oracle's; the other 125 are further bench windows. Tokens per the free-running statement, prosody
    tags equal, packets byte-identical wherever the text agrees."""
    from janus_b200.pipeline import SemanticEncoder, join_segments, split_text_segments
    from janus_b200.protocol import JanusMode, JanusPacket
    from janus_b200.synth import make_clips
    from oracle.prosody_oracle import ProsodyOracle
    k, B = 3, 128
    dims, spec, W, clips, enc_ref, ref_tok, ref_lp, margins = _oracle("small", k)
    enc = SemanticEncoder("small", batch=B, max_new_tokens=N_NEW, weights=W, lane_batch=B, n_lanes=1)
    assert enc.session.latent
    rest = make_clips(B - k, 30.0, 16000, seed=1234, first=k).numpy()
    bufs = list(clips) + [rest[i] for i in range(B - k)]
    lengths = tuple(len(b) for b in bufs)
    pcm = torch.from_numpy(np.concatenate(bufs)).cuda()
    res = enc.encode_device(pcm, lengths).synchronize()
    tok = res.tokens.cpu()
    exact, _ = check_free_running(tok[:k], ref_tok, margins, MARGIN_TOL["small"], "engine lane-128")
    texts, pkts = enc.to_packets(tok[:k].numpy(), res.tags[:k].cpu().numpy(), timestamps=[1.0] * k)
    n_pkt_equal = 0
    for i in range(k):
        otag = ProsodyOracle(sample_rate=16000).analyze_buffer(bufs[i], python_hop_loop=False)
        ref_text = join_segments(split_text_segments(ref_tok[i].tolist(), spec))
        ref_pkt = JanusPacket(ref_text, JanusMode.SEMANTIC_VOICE, otag, None, 1.0).serialize()
        got = JanusPacket.deserialize(pkts[i])
        assert got.prosody == otag, (i, got.prosody, otag)
        if texts[i] == ref_text:
            assert pkts[i] == ref_pkt
            n_pkt_equal += 1
    print(f"engine lane-128: {exact}/{k} windows bit-exact, {n_pkt_equal}/{k} packets byte-identical, tags equal for all")
    # determinism of the route: a second call reproduces the first bit for bit
    res2 = enc.encode_device(pcm, lengths).synchronize()
    assert torch.equal(res2.tokens.cpu(), tok)


def test_large_v3_full_depth_encoder_greedy_and_beam5_spot_check():
    """BASELINE configs[3] at FULL depth: Whisper-large-v3 (128 mels, d = 1280, 20 heads, 32 + 32 layers, 1.55 B
    random-init parameters), one 12 s window from audio: encoder states against the fp32 oracle, greedy tokens per the
    free-running statement, and the beam-5 best hypothesis where the oracle's result is stable under +-0.03 logit noise.
    Short (6 new tokens): the fp32 CPU oracle reads 6 GB of weights per decode step."""
    from janus_b200.synth import make_clips
    from janus_b200.whisper.config import PRESETS, tokenizer_spec
    from janus_b200.whisper.engine import DecodeSession, WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = PRESETS["large-v3"]
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=0, profile="confident")
    clip = make_clips(1, 12.0, 16000, seed=1234).numpy()[0]
    n_new = 6
    with torch.no_grad():
        mel = torch.from_numpy(log_mel_window(clip, dims.n_mels))[None]
        enc_ref = R.encode(W, dims, mel)
        g_tok, g_lp, margins = R.greedy_decode(W, dims, spec, enc_ref, n_new, return_margins=True)
        b_tok, b_lp = R.beam_decode(W, dims, spec, enc_ref, n_new, beam_size=5)
        stable = R.beam_stable_windows(W, dims, spec, enc_ref, n_new, b_tok, beam_size=5, trials=2)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    enc = _gpu_encode(eng, [clip])
    err = (enc.float().cpu() - enc_ref).abs()
    print(f"large-v3 encoder (32 layers) vs fp32 oracle: mean |err| {err.mean():.4g} max {err.max():.4g}")
    assert err.mean() <= 2e-2 and err.max() <= 0.5
    ses = DecodeSession(eng, batch=1, max_new_tokens=n_new)
    toks, _ = ses.decode(enc)
    torch.cuda.synchronize()
    check_free_running(toks, g_tok, margins, 0.25, "large-v3 greedy")
    bs = DecodeSession(eng, batch=1, max_new_tokens=n_new, beam=5)
    bt, blp = bs.decode(enc)
    torch.cuda.synchronize()
    print(f"large-v3 beam-5: oracle result stable under logit noise: {stable[0]}; tokens {bt[0].tolist()} vs {b_tok[0].tolist()}")
    if stable[0]:
        assert bt[0].cpu().long().tolist() == b_tok[0].tolist()
        assert abs(float(blp[0]) - float(b_lp[0])) <= 0.3
