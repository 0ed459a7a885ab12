"""K3 beam search on a real B200 (through the C ABI: janus_whisper_decode_beam_step / janus_whisper_beam_finalize)
vs the fp32 CPU oracle's beam_decode (oracle/whisper_ref.py).

Stated bars:
  * beam = 1 is the greedy decoder: tokens identical to the greedy session of the same engine;
  * "confident" weights (margins >> bf16 noise): the best hypothesis of every window equals the oracle's wherever
    the oracle's best / second-best normalised scores differ by more than SCORE_MARGIN, its sum of log-probs is
    within LP_TOL, and the list of finished hypotheses (tokens, order, lengths) equals the oracle's wherever the
    scores that ordered it are SCORE_MARGIN apart -- for windows whose oracle result is STABLE, i.e. reproduced under
    +-0.03 logit perturbations (oracle.whisper_ref.beam_stable_windows): a beam search takes 30 discrete decisions per
    step and a near-tie at any of them may resolve differently under bf16 rounding;
  * the grouped latent cross-attention (beams sharing one window of encoder states) is bit-identical to the
    ungrouped kernel on replicated encoder states.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
SCORE_MARGIN = 0.05
LP_TOL = 0.15


def _setup(eot_scale=None, d=128, heads=2, layers=2, n_clips=3, seed=3):
    from janus_b200.synth import make_clips
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.engine import WhisperEngine
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = mini_dims(d=d, heads=heads, layers=layers)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=seed, profile="confident")
    if eot_scale is not None:
        # make <|endoftext|> a strong runner-up of the most frequent token so that hypotheses finish at different
        # lengths, finished lists fill up and windows complete early
        with torch.no_grad():
            clips0 = make_clips(1, 30.0, 16000, seed=77).numpy()
            mel0 = torch.from_numpy(log_mel_window(clips0[0], dims.n_mels))[None]
            g, _ = R.greedy_decode(W, dims, spec, R.encode(W, dims, mel0), 4)
        frequent = int(g[0, 1])
        E = W["dec.tok_emb"].clone()
        E[spec.eot] = (eot_scale * E[frequent]).to(torch.bfloat16).float()
        W = dict(W)
        W["dec.tok_emb"] = E
    clips = make_clips(n_clips, 30.0, 16000, seed=77).numpy()
    mel = torch.stack([torch.from_numpy(log_mel_window(c, dims.n_mels)) for c in clips])
    with torch.no_grad():
        enc_ref = R.encode(W, dims, mel)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    return dims, spec, W, enc_ref, eng, R


@pytest.mark.parametrize("latent", [False, True])
def test_beam1_is_greedy(latent):
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, enc_ref, eng, R = _setup()
    enc = enc_ref.to(torch.bfloat16).cuda().contiguous()
    n_new = 24
    g = DecodeSession(eng, batch=enc.shape[0], max_new_tokens=n_new, latent=latent)
    gt, glp = g.decode(enc)
    b = DecodeSession(eng, batch=enc.shape[0], max_new_tokens=n_new, latent=latent, beam=1)
    bt, blp = b.decode(enc)
    torch.cuda.synchronize()
    assert torch.equal(gt.cpu(), bt.cpu())
    # latent: the same kernels run in both sessions; K/V form: the beam session's cross-attention is the tensor-core
    # kernel (p rounded to bf16 for the second MMA), the greedy session's the fp32 FMA kernel
    assert torch.allclose(glp.cpu(), blp.cpu(), atol=1e-3 if latent else LP_TOL)


def _norm(score, length):
    return score / max(length, 1)


@pytest.mark.parametrize("latent,eot_scale,use_graph,d,heads", [
    (False, None, True, 128, 2), (True, None, True, 128, 2), (False, 0.9, False, 128, 2), (True, 0.9, True, 128, 2),
    (False, 0.9, True, 384, 6), (True, 0.9, True, 384, 6)])
def test_beam5_vs_oracle(latent, eot_scale, use_graph, d, heads):
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, enc_ref, eng, R = _setup(eot_scale, d=d, heads=heads)
    n_new, K = 16, 5
    with torch.no_grad():
        ref_tok, ref_lp, pools, finished = R.beam_decode(W, dims, spec, enc_ref, n_new, beam_size=K, return_all=True)
        stable = R.beam_stable_windows(W, dims, spec, enc_ref, n_new, ref_tok, beam_size=K)
    enc = enc_ref.to(torch.bfloat16).cuda().contiguous()
    ses = DecodeSession(eng, batch=enc.shape[0], max_new_tokens=n_new, latent=latent, beam=K, use_graph=use_graph,
                        steps_per_graph=4)
    assert ses.latent == latent
    for rep in range(2):   # the second decode reuses the session (reset of cum / indirection / finished lists)
        tok, lp = ses.decode(enc)
        torch.cuda.synchronize()
        tok, lp = tok.cpu().long(), lp.cpu()
        n_fin = ses.n_fin.cpu().tolist()
        checked = 0
        for b in range(enc.shape[0]):
            pool = pools[b]
            ranked = sorted((_norm(p[1], p[2]) for p in pool), reverse=True)
            if stable[b] and (len(ranked) < 2 or ranked[0] - ranked[1] > SCORE_MARGIN):
                assert tok[b].tolist() == ref_tok[b].tolist(), (b, tok[b].tolist(), ref_tok[b].tolist())
                assert abs(float(lp[b]) - float(ref_lp[b])) <= LP_TOL
                checked += 1
            # finished hypotheses: the oracle's pool lists them first, in the order they finished
            fin_ref = finished[b]
            if eot_scale is not None:
                fs = ses.fin_score[b].cpu().tolist()
                fl = ses.fin_len[b].cpu().tolist()
                ft = ses.fin_tokens[b].cpu().tolist()
                scores = [p[1] for p in fin_ref]
                distinct = all(abs(scores[i] - scores[j]) > SCORE_MARGIN for i in range(len(scores))
                               for j in range(i + 1, len(scores)))
                if distinct and stable[b] and len(fin_ref) == n_fin[b]:
                    for j, p in enumerate(fin_ref):
                        assert fl[j] == p[2], (b, j, fl, [q[2] for q in fin_ref])
                        assert ft[j][ses.n_prompt:ses.n_prompt + p[2]] == p[0], (b, j)
                        assert ft[j][ses.n_prompt + p[2]] == spec.eot
                        assert abs(fs[j] - p[1]) <= LP_TOL
        assert checked >= 1, "no window had a decisive best hypothesis: the test checked nothing"
        if eot_scale is not None:
            assert sum(n_fin) > 0, "the eot-biased weights were meant to finish hypotheses"
            done = ses.stream_done.cpu().tolist()
            assert [int(n >= K) for n in n_fin] == done


@pytest.mark.parametrize("layout", [0, 1])
@pytest.mark.parametrize("n_streams,beam,T,H", [(3, 5, 1500, 12), (9, 2, 700, 6), (4, 8, 1500, 20), (40, 5, 333, 12),
                                                (2, 3, 70, 2)])
def test_cross_attn_beam_kernel_equals_per_row_kernel(n_streams, beam, T, H, layout):
    """The K/V cross-attention kernels that serve all beams of a stream from one pass over K / V (layout 0: packed
    FMAs on the greedy layouts; layout 1: mma.sync on K row-major / V transposed) against the greedy kernel on
    replicated K / V. Bar: 4e-3 of the output scale (layout 1 rounds p to bf16 for the second MMA)."""
    from janus_b200 import _abi
    g = torch.Generator(device="cuda").manual_seed(5)
    d, kv_ld = 64 * H, -(-T // 64) * 64
    rows = n_streams * beam
    q = torch.randn(rows, d, device="cuda", generator=g).bfloat16()
    kT = torch.randn(n_streams, H, 64, kv_ld, device="cuda", generator=g).bfloat16()
    v = torch.randn(n_streams, H, T, 64, device="cuda", generator=g).bfloat16()
    o1 = torch.full((rows, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    o2 = torch.full((rows, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    if layout == 1:
        k_in = kT[..., :T].transpose(-1, -2).contiguous()                      # [n, H, T, 64]
        v_in = torch.full((n_streams, H, 64, kv_ld), float("nan"), device="cuda", dtype=torch.bfloat16)
        v_in[..., :T] = v.transpose(-1, -2)                                     # pad keys stay NaN: must be masked
    else:
        k_in, v_in = kT, v
    _abi.call("janus_dec_cross_attn_beam_bf16", q.data_ptr(), k_in.data_ptr(), v_in.data_ptr(), o1.data_ptr(), n_streams,
              beam, H, T, kv_ld, layout, st)
    kr, vr = kT.repeat_interleave(beam, 0).contiguous(), v.repeat_interleave(beam, 0).contiguous()
    _abi.call("janus_dec_cross_attn_bf16", q.data_ptr(), kr.data_ptr(), vr.data_ptr(), o2.data_ptr(), rows, H, T, kv_ld, st)
    torch.cuda.synchronize()
    assert torch.isfinite(o1.float()).all()
    err = (o1.float() - o2.float()).abs().max().item()
    assert err <= 4e-3 * max(1.0, o2.float().abs().max().item()), err


@pytest.mark.parametrize("n_streams,group,T,d,H", [(7, 5, 1500, 768, 12), (40, 5, 300, 768, 12), (6, 4, 200, 384, 6),
                                                   (31, 5, 500, 1024, 16)])
def test_xlat_grouped_is_bit_identical(n_streams, group, T, d, H):
    from janus_b200 import _abi
    g = torch.Generator(device="cuda").manual_seed(11)
    n = n_streams * group
    enc = torch.randn(n_streams, T, d, device="cuda", generator=g).bfloat16()
    qlat = torch.zeros(n, 16, d, device="cuda", dtype=torch.bfloat16)
    qlat[:, :H] = (torch.randn(n, H, d, device="cuda", generator=g) * (8.0 / d ** 0.5)).bfloat16()
    z1 = torch.full((n, H, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    z2 = torch.full((n, H, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    _abi.call("janus_dec_xlat_grouped_bf16", qlat.data_ptr(), enc.data_ptr(), z1.data_ptr(), n, group, T, d, H, st)
    rep = enc.repeat_interleave(group, 0).contiguous()
    _abi.call("janus_dec_xlat_bf16", qlat.data_ptr(), rep.data_ptr(), z2.data_ptr(), n, T, d, H, st)
    torch.cuda.synchronize()
    assert torch.isfinite(z1.float()).all()
    assert torch.equal(z1, z2)


@pytest.mark.parametrize("latent", [False, True])
def test_beam5_many_windows_replica_property(latent):
    """Size-independent property at a lane-sized batch (136 windows = 680 decode rows, tcgen05 GEMM path, latent form
    at >= 128 windows): replicas of a window, scattered over the batch, must produce identical hypotheses, finished
    lists and scores -- whatever their neighbours do to the shared buffers, the indirection and the reorder."""
    from janus_b200.whisper.engine import DecodeSession
    dims, spec, W, enc_ref, eng, R = _setup(0.9, d=384, heads=6, layers=2, n_clips=8)
    n_rep = 17
    idx = torch.arange(8).repeat(n_rep)[torch.randperm(8 * n_rep, generator=torch.Generator().manual_seed(1))]
    enc = enc_ref.to(torch.bfloat16)[idx].cuda().contiguous()
    ses = DecodeSession(eng, batch=enc.shape[0], max_new_tokens=12, latent=latent, beam=5, steps_per_graph=5)
    assert ses.latent == latent
    tok, lp = ses.decode(enc)
    torch.cuda.synchronize()
    tok, lp, nf = tok.cpu(), lp.cpu(), ses.n_fin.cpu()
    fin_t, fin_s = ses.fin_tokens.cpu(), ses.fin_score.cpu()
    for w in range(8):
        rows = (idx == w).nonzero().flatten().tolist()
        r0 = rows[0]
        for r in rows[1:]:
            assert torch.equal(tok[r], tok[r0]) and float(lp[r]) == float(lp[r0]), (w, r)
            assert int(nf[r]) == int(nf[r0])
            n = int(nf[r0])
            assert torch.equal(fin_t[r, :n], fin_t[r0, :n]) and torch.equal(fin_s[r, :n], fin_s[r0, :n])
    assert int(nf.sum()) > 0


@pytest.mark.parametrize("case", ["random", "eot_biased"])
@pytest.mark.parametrize("latent", [False, True])
def test_decode_matches_committed_golden_vectors(case, latent):
    """Device greedy and beam-5 decodes against the committed golden vectors (tests/golden/decode_golden.json): greedy
    tokens of every window; beam tokens / sums of log-probs where the golden result is decisive; finished lists
    where their scores are SCORE_MARGIN apart."""
    import json
    import os
    import sys
    from janus_b200.whisper.engine import DecodeSession, WhisperEngine
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, here)
    try:
        import make_golden
    finally:
        sys.path.remove(here)
    gold = json.load(open(os.path.join(here, "decode_golden.json")))
    g = gold["cases"][case]
    dims, spec, W, enc_ref, R = make_golden.decode_setup(None if case == "random" else 0.9)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    enc = enc_ref.to(torch.bfloat16).cuda().contiguous()
    n_new, K = gold["n_new"], gold["beam"]
    gs = DecodeSession(eng, batch=enc.shape[0], max_new_tokens=n_new, latent=latent)
    gt, glp = gs.decode(enc)
    bs = DecodeSession(eng, batch=enc.shape[0], max_new_tokens=n_new, latent=latent, beam=K)
    bt, blp = bs.decode(enc)
    torch.cuda.synchronize()
    gt, bt = gt.cpu().tolist(), bt.cpu().tolist()
    # greedy: tokens up to (and including) the first eot are the contract; the oracle pads with eot afterwards
    for b in range(enc.shape[0]):
        ref = g["greedy_tokens"][b]
        n = ref.index(spec.eot) + 1 if spec.eot in ref else len(ref)
        assert gt[b][:n] == ref[:n], (b, gt[b], ref)
        assert abs(float(glp[b]) - g["greedy_sum_logprob"][b]) <= LP_TOL
        if g["beam_decisive"][b]:
            assert bt[b] == g["beam_tokens"][b], (b, bt[b], g["beam_tokens"][b])
            assert abs(float(blp[b]) - g["beam_sum_logprob"][b]) <= LP_TOL
        fin = g["finished"][b]
        scores = [f["score"] for f in fin]
        if g["beam_decisive"][b] and all(abs(scores[i] - scores[j]) > SCORE_MARGIN for i in range(len(scores))
                                         for j in range(i + 1, len(scores))):
            assert int(bs.n_fin[b]) == len(fin)
            ft, fl = bs.fin_tokens[b].cpu().tolist(), bs.fin_len[b].cpu().tolist()
            for j, f in enumerate(fin):
                assert fl[j] == f["len"] and ft[j][bs.n_prompt:bs.n_prompt + f["len"]] == f["tokens"]
    assert any(g["beam_decisive"])
