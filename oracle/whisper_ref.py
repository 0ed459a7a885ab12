"""CPU oracle for Whisper STT (encoder, KV-cache decoder, greedy decode) -- TEST INFRASTRUCTURE ONLY.

fp32 PyTorch restatement of what the reference reaches through
`WhisperModel(...).transcribe(audio_16k, beam_size=1, language='en')`
(/root/reference/backend/services/transcriber.py:23-27,53-57). The arithmetic lives in the
un-vendored third-party packages faster-whisper (>=1.0.0, requirements.txt:2) -> CTranslate2; what is
restated here is the published OpenAI Whisper architecture those packages execute and the decode-time
logit rules of faster-whisper's `generate_with_fallback` at temperature 0 / beam_size 1:
suppress_blank, suppress_tokens=[-1] (non-speech set + specials), timestamp rules with
max_initial_timestamp_index=50, max_length 448.

PARITY UNPINNED against faster-whisper/CTranslate2 (not installable offline; the reference's tests mock
the model). Cross-checked against the independent HF `transformers` Whisper implementation in
tests/test_oracle_stt.py (same weights, fp32): encoder states and teacher-forced logits; apply_logit_rules against HF's
own Whisper logits processors (identical forbidden sets); greedy_decode against HF generate with those processors (token
for token); beam_decode against HF's generic beam search with this module's rules as a renormalising LogitsProcessor.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from janus_b200.whisper.config import TokenizerSpec, WhisperDims


def _ln(x, W, p):
    return F.layer_norm(x, (x.shape[-1],), W[p + ".w"], W[p + ".b"], eps=1e-5)


def _lin(x, W, p):
    return F.linear(x, W[p + ".w"], W.get(p + ".b"))


def _mha(xq, xkv, W, p, n_head, mask=None, kv=None):
    """Whisper MultiHeadAttention: scale d_h**-0.25 on both q and k. kv = precomputed (k, v)."""
    q = _lin(xq, W, p + ".q")
    if kv is None:
        k, v = _lin(xkv, W, p + ".k"), _lin(xkv, W, p + ".v")
    else:
        k, v = kv
    B, Tq, D = q.shape
    dh = D // n_head
    scale = dh ** -0.25
    qh = (q.view(B, Tq, n_head, dh).permute(0, 2, 1, 3)) * scale
    kh = (k.view(B, -1, n_head, dh).permute(0, 2, 3, 1)) * scale
    vh = v.view(B, -1, n_head, dh).permute(0, 2, 1, 3)
    s = qh @ kh
    if mask is not None:
        s = s + mask
    w = torch.softmax(s.float(), dim=-1)
    o = (w @ vh).permute(0, 2, 1, 3).reshape(B, Tq, D)
    return _lin(o, W, p + ".o"), (k, v)


def encode(W, dims: WhisperDims, mel: torch.Tensor, taps: dict | None = None) -> torch.Tensor:
    """mel: [B, n_mels, 2*n_audio_ctx] fp32 -> [B, n_audio_ctx, d]."""
    x = F.gelu(F.conv1d(mel, W["enc.conv1.w"], W["enc.conv1.b"], padding=1))
    if taps is not None:
        taps["conv1"] = x.permute(0, 2, 1)
    x = F.gelu(F.conv1d(x, W["enc.conv2.w"], W["enc.conv2.b"], stride=2, padding=1))
    x = x.permute(0, 2, 1) + W["enc.pos"]
    if taps is not None:
        taps["conv2"] = x
    for i in range(dims.n_audio_layer):
        p = f"enc.l{i}"
        h = _ln(x, W, p + ".attn_ln")
        a, _ = _mha(h, h, W, p + ".attn", dims.n_audio_head)
        x = x + a
        h = _lin(F.gelu(_lin(_ln(x, W, p + ".mlp_ln"), W, p + ".mlp.fc1")), W, p + ".mlp.fc2")
        x = x + h
        if taps is not None:
            taps[f"l{i}"] = x
    return _ln(x, W, "enc.ln_post")


class DecoderState:
    """Self-attention KV cache + per-layer cross K/V (computed once per window)."""

    def __init__(self, W, dims: WhisperDims, enc_out: torch.Tensor):
        self.W, self.dims = W, dims
        self.cross = []
        for i in range(dims.n_text_layer):
            p = f"dec.l{i}.cross"
            self.cross.append((_lin(enc_out, W, p + ".k"), _lin(enc_out, W, p + ".v")))
        self.self_kv = [None] * dims.n_text_layer
        self.offset = 0

    def step(self, tokens: torch.Tensor) -> torch.Tensor:
        """tokens: [B, T_new] -> logits [B, T_new, V] (fp32), appending to the cache."""
        W, dims = self.W, self.dims
        T = tokens.shape[1]
        x = W["dec.tok_emb"][tokens] + W["dec.pos"][self.offset:self.offset + T]
        total = self.offset + T
        mask = torch.full((T, total), float("-inf")).triu_(self.offset + 1)
        for i in range(dims.n_text_layer):
            p = f"dec.l{i}"
            h = _ln(x, W, p + ".attn_ln")
            k_new, v_new = _lin(h, W, p + ".attn.k"), _lin(h, W, p + ".attn.v")
            if self.self_kv[i] is not None:
                k_all = torch.cat([self.self_kv[i][0], k_new], 1)
                v_all = torch.cat([self.self_kv[i][1], v_new], 1)
            else:
                k_all, v_all = k_new, v_new
            self.self_kv[i] = (k_all, v_all)
            a, _ = _mha(h, None, W, p + ".attn", dims.n_text_head, mask=mask, kv=(k_all, v_all))
            x = x + a
            c, _ = _mha(_ln(x, W, p + ".cross_ln"), None, W, p + ".cross", dims.n_text_head, kv=self.cross[i])
            x = x + c
            x = x + _lin(F.gelu(_lin(_ln(x, W, p + ".mlp_ln"), W, p + ".mlp.fc1")), W, p + ".mlp.fc2")
        self.offset = total
        x = _ln(x, W, "dec.ln")
        return (x @ W["dec.tok_emb"].T).float()


def apply_logit_rules(logits: torch.Tensor, generated: list[int], spec: TokenizerSpec,
                      suppress_mask: torch.Tensor, return_gap: bool = False):
    """One sequence. `generated` = tokens sampled so far (after the sot prompt). Returns filtered
    logits (-inf = forbidden). Order follows faster-whisper / openai-whisper: SuppressBlank,
    SuppressTokens, ApplyTimestampRules."""
    lg = logits.clone()
    ninf = float("-inf")
    tb = spec.timestamp_begin
    if len(generated) == 0:                       # SuppressBlank at sample_begin
        lg[spec.blank] = ninf
        lg[spec.eot] = ninf
    lg[suppress_mask] = ninf                       # SuppressTokens (+ <|notimestamps|>)
    last_ts = len(generated) >= 1 and generated[-1] >= tb
    penult_ts = len(generated) < 2 or generated[-2] >= tb
    if last_ts:
        if penult_ts:
            lg[tb:] = ninf                         # must be non-timestamp
        else:
            lg[:spec.eot] = ninf                   # cannot be a normal text token
    ts = [t for t in generated if t >= tb]
    if ts:
        ts_last = ts[-1] if (last_ts and not penult_ts) else ts[-1] + 1
        lg[tb:ts_last] = ninf                      # timestamps never decrease
    if len(generated) == 0:
        lg[:tb] = ninf                             # first token must be a timestamp ...
        lg[tb + spec.max_initial_timestamp_index + 1:] = ninf   # ... no later than 1.0 s
    lp = torch.log_softmax(lg.float(), dim=-1)
    ts_lp = torch.logsumexp(lp[tb:], dim=-1)
    max_text = lp[:tb].max()
    if ts_lp > max_text:
        lg[:tb] = ninf
    if return_gap:
        # how close the rule's own decision was: |log sum p(timestamps) - max log p(text)| when both categories are
        # open. A near-tie HERE flips the sampled token between a timestamp and a text token although the top-1 / top-2
        # margin inside either category may be large, so the decision margin of a greedy step is the min of the two.
        both = bool(torch.isfinite(ts_lp)) and bool(torch.isfinite(max_text))
        return lg, (float((ts_lp - max_text).abs()) if both else float("inf"))
    return lg


def apply_logit_rules_batch(logits: torch.Tensor, generated: list[list[int]], spec: TokenizerSpec,
                            suppress_mask: torch.Tensor, active=None, return_gap: bool = False):
    """apply_logit_rules() for B sequences at once (same rules, same order, as column-range masks), so that the CPU
    baseline is not dominated by a per-sequence Python loop. Checked against the scalar function on random
    histories in tests/test_oracle_stt.py. Rows with active[b] == False are returned unfiltered."""
    B, V = logits.shape
    tb, eot = spec.timestamp_begin, spec.eot
    ninf = float("-inf")
    lg = logits.clone()
    col = torch.arange(V)[None, :]
    n = torch.tensor([len(g) for g in generated])
    act = torch.ones(B, dtype=torch.bool) if active is None else torch.as_tensor(active, dtype=torch.bool)
    first = (n == 0) & act
    last_ts = torch.tensor([len(g) >= 1 and g[-1] >= tb for g in generated]) & act
    penult_ts = torch.tensor([len(g) < 2 or g[-2] >= tb for g in generated])
    # the most recent timestamp token so far (-1: none)
    last_stamp = torch.tensor([next((t for t in reversed(g) if t >= tb), -1) for g in generated])
    forbid = suppress_mask[None, :] & act[:, None]
    forbid = forbid | (first[:, None] & ((col == spec.blank) | (col == eot)))
    forbid = forbid | ((last_ts & penult_ts)[:, None] & (col >= tb))
    forbid = forbid | ((last_ts & ~penult_ts)[:, None] & (col < eot))
    floor = torch.where(last_ts & ~penult_ts, last_stamp, last_stamp + 1)
    forbid = forbid | (((last_stamp >= 0) & act)[:, None] & (col >= tb) & (col < floor[:, None]))
    forbid = forbid | (first[:, None] & ((col < tb) | (col > tb + spec.max_initial_timestamp_index)))
    lg[forbid] = ninf
    lp = torch.log_softmax(lg.float(), dim=-1)
    ts_lp = torch.logsumexp(lp[:, tb:], dim=-1)
    max_text = lp[:, :tb].max(dim=-1).values
    lg[((ts_lp > max_text) & act)[:, None] & (col < tb)] = ninf
    if return_gap:
        gap = (ts_lp - max_text).abs()
        gap[~(torch.isfinite(ts_lp) & torch.isfinite(max_text))] = float("inf")
        return lg, gap
    return lg


def suppress_mask_for(spec: TokenizerSpec) -> torch.Tensor:
    m = torch.zeros(spec.n_vocab, dtype=torch.bool)
    m[list(spec.suppress_ids)] = True
    return m


def greedy_decode(W, dims: WhisperDims, spec: TokenizerSpec, enc_out: torch.Tensor, max_new_tokens: int,
                  stop_at_eot: bool = True, return_margins: bool = False, vectorized: bool = False,
                  prompt: list | None = None, return_no_speech: bool = False):
    """Batched greedy decode at temperature 0. Returns tokens [B, max_new_tokens] (eot-padded after
    the first eot), sum of log-probs [B], and optionally the DECISION MARGIN of every step: the smaller of the
    top-1 / top-2 gap of the rule-filtered logits and the gap of the timestamps-vs-text rule decision (in
    log-prob units) -- the amount of logit error that could change the sampled token."""
    B = enc_out.shape[0]
    st = DecoderState(W, dims, enc_out)
    sup = suppress_mask_for(spec)
    # `prompt`: the forced prefix of every row (default: the sot sequence); with previous-text conditioning it is
    # [<|startofprev|>, previous tokens ..., <|startoftranscript|>, ...]. no_speech_prob = softmax of the logits at the
    # <|startoftranscript|> position (faster-whisper return_no_speech_prob / openai-whisper DecodingTask._main_loop)
    plist = list(prompt) if prompt is not None else list(spec.sot_sequence)
    all_logits = st.step(torch.tensor([plist] * B, dtype=torch.long))
    logits = all_logits[:, -1]
    sot_index = max(i for i, t in enumerate(plist) if t == spec.sot) if spec.sot in plist else 0
    no_speech = torch.softmax(all_logits[:, sot_index].float(), dim=-1)[:, spec.no_speech]
    gen = [[] for _ in range(B)]
    done = [False] * B
    out = torch.full((B, max_new_tokens), spec.eot, dtype=torch.long)
    sum_lp = torch.zeros(B)
    margins = torch.full((B, max_new_tokens), float("inf"))
    n_steps = min(max_new_tokens, spec.max_length - len(plist))
    for t in range(n_steps):
        nxt = torch.full((B,), spec.eot, dtype=torch.long)
        if vectorized:   # the same rules for all rows at once (apply_logit_rules_batch)
            lg_all, gap_all = apply_logit_rules_batch(logits, gen, spec, sup, active=[not d_ for d_ in done],
                                                      return_gap=True)
        for b in range(B):
            if done[b]:
                continue
            if vectorized:
                lg, gap = lg_all[b], float(gap_all[b])
            else:
                lg, gap = apply_logit_rules(logits[b], gen[b], spec, sup, return_gap=True)
            lp = torch.log_softmax(lg, dim=-1)
            top2 = torch.topk(lg, 2)
            tok = int(top2.indices[0])
            # decision margin of this step: top-1 vs top-2 of the filtered logits, or the gap of the
            # "timestamps outweigh text" rule itself, whichever is smaller
            margins[b, t] = min(float(top2.values[0] - top2.values[1]), gap)
            nxt[b] = tok
            out[b, t] = tok
            sum_lp[b] += lp[tok]
            gen[b].append(tok)
            if tok == spec.eot and stop_at_eot:
                done[b] = True
        if all(done) or t == n_steps - 1:
            break
        logits = st.step(nxt[:, None])[:, -1]
    extra = ((margins,) if return_margins else ()) + ((no_speech,) if return_no_speech else ())
    return (out, sum_lp) + extra


# ----------------------------------------------------------------------------------------------
# beam search (BASELINE configs[3]: beam 5; the reference's own call uses beam_size=1, transcriber.py:55)
# ----------------------------------------------------------------------------------------------
def beam_step_select(cand: list[tuple[float, int, int]], beam_size: int, eot: int):
    """One stream, one step of the beam bookkeeping. `cand` = (score, beam, token) for the top
    beam_size + 1 tokens of every live beam. Walk the candidates by descending score (ties: beam, then
    token ascending; a score of -inf is no candidate): eot candidates finish their hypothesis, the first
    `beam_size` others become the next beams; slots left over are dead beams (score -inf, their own history,
    token eot). Returns (next beams [(score, parent, token)], newly finished [(score, parent)])."""
    order = sorted(cand, key=lambda c: (-c[0], c[1], c[2]))
    nxt, fin = [], []
    for s, k, t in order:
        if s == float("-inf"):
            break
        if t == eot:
            fin.append((s, k))
        else:
            nxt.append((s, k, t))
            if len(nxt) == beam_size:
                break
    while len(nxt) < beam_size:
        nxt.append((float("-inf"), len(nxt), eot))
    return nxt, fin


def beam_decode(W, dims: WhisperDims, spec: TokenizerSpec, enc_out: torch.Tensor, max_new_tokens: int,
                beam_size: int = 5, patience: float = 1.0, return_all: bool = False, logit_noise: float = 0.0,
                noise_seed: int = 0):
    """Beam search at temperature 0 as published in OpenAI Whisper's `BeamSearchDecoder.update/finalize` +
    `MaximumLikelihoodRanker(length_penalty=None)` -- the algorithm faster-whisper / CTranslate2 run for
    `beam_size=5, patience=1, length_penalty=1` (PARITY UNPINNED against CTranslate2, see the module header):

    * per live beam the rule-filtered log-softmax; candidates = its top beam_size+1 tokens, score = cumulative
      log-prob + token log-prob; the beams of a stream are identical before the first sampled token, so only
      beam 0 is live there (cumulative score -inf for the others: same effect as the dict de-duplication);
    * beam_step_select(); newly finished hypotheses join the stream's finished list in descending score while
      it holds fewer than max_candidates = round(beam_size * patience); a stream is complete when it is full;
    * at the end a stream with fewer than beam_size finished hypotheses adds its live beams by descending
      cumulative score; the result is the hypothesis with the largest score / n_tokens (tokens before eot).

    Returns tokens [B, max_new_tokens] (eot-padded), sum_logprob [B] of the chosen hypothesis; with
    return_all also the per-stream ranking pools and finished lists [(tokens, score, length)]."""
    B, K = enc_out.shape[0], beam_size
    max_cand = int(round(beam_size * patience))
    eot = spec.eot
    st = DecoderState(W, dims, enc_out.repeat_interleave(K, 0))
    sup = suppress_mask_for(spec)
    prompt = torch.tensor([list(spec.sot_sequence)] * (B * K), dtype=torch.long)
    ng = torch.Generator().manual_seed(noise_seed)

    def perturb(lg):   # stability probe (beam_stable_windows): uniform logit noise of the size of bf16 rounding
        return lg + logit_noise * (2 * torch.rand(lg.shape, generator=ng) - 1) if logit_noise else lg
    logits = perturb(st.step(prompt)[:, -1])
    gen = [[[] for _ in range(K)] for _ in range(B)]
    cum = [[0.0] + [float("-inf")] * (K - 1) for _ in range(B)]
    finished = [[] for _ in range(B)]     # (tokens, score, length)
    n_steps = min(max_new_tokens, spec.max_length - len(spec.sot_sequence))
    f32 = lambda v: float(torch.as_tensor(v, dtype=torch.float32))
    for t in range(n_steps):
        src = torch.arange(B * K)
        nxt_tok = torch.full((B * K,), eot, dtype=torch.long)
        for b in range(B):
            if len(finished[b]) >= max_cand:
                continue
            cand = []
            for k in range(K):
                lg = apply_logit_rules(logits[b * K + k], gen[b][k], spec, sup)
                lp = torch.log_softmax(lg, dim=-1)
                v, i = torch.sort(lp, descending=True, stable=True)
                for r in range(K + 1):
                    cand.append((f32(torch.tensor(cum[b][k], dtype=torch.float32) + v[r]), k, int(i[r])))
            nxt, fin = beam_step_select(cand, K, eot)
            for s, k in fin:
                if len(finished[b]) < max_cand:
                    finished[b].append((list(gen[b][k]), s, len(gen[b][k])))
            new_gen, new_cum = [], []
            for j, (s, k, tok) in enumerate(nxt):
                new_gen.append(gen[b][k] + [tok])
                new_cum.append(s)
                src[b * K + j] = b * K + k
                nxt_tok[b * K + j] = tok
            gen[b], cum[b] = new_gen, new_cum
        if all(len(f) >= max_cand for f in finished) or t == n_steps - 1:
            break
        for i in range(dims.n_text_layer):
            st.self_kv[i] = (st.self_kv[i][0][src], st.self_kv[i][1][src])
        logits = perturb(st.step(nxt_tok[:, None])[:, -1])
    out = torch.full((B, max_new_tokens), eot, dtype=torch.long)
    sum_lp = torch.zeros(B)
    pools = []
    for b in range(B):
        pool = list(finished[b])
        if len(pool) < K:
            for k in sorted(range(K), key=lambda k: (-cum[b][k], k)):
                pool.append((list(gen[b][k]), cum[b][k], len(gen[b][k])))
                if len(pool) >= K:
                    break
        best = max(range(len(pool)), key=lambda j: (f32(pool[j][1]) / max(pool[j][2], 1), -j))
        toks = pool[best][0][:max_new_tokens]
        out[b, :len(toks)] = torch.tensor(toks, dtype=torch.long)
        sum_lp[b] = pool[best][1]
        pools.append(pool)
    if return_all:
        return out, sum_lp, pools, finished
    return out, sum_lp


def beam_stable_windows(W, dims: WhisperDims, spec: TokenizerSpec, enc_out: torch.Tensor, max_new_tokens: int,
                        ref_tokens: torch.Tensor, beam_size: int = 5, noise: float = 0.03, trials: int = 3) -> list[bool]:
    """Which windows have a beam-search result that a bf16 implementation can be held to: a beam search takes
    beam_size x (beam_size + 1) discrete decisions per step, so a near-tie at ANY step may legitimately resolve
    differently under bf16 rounding and change which hypotheses reach the pool. A window is `stable` when the
    unperturbed result `ref_tokens` is reproduced under `trials` independent uniform logit perturbations of +-`noise`
    (the size of the bf16 logit error measured on the GPU)."""
    stable = [True] * enc_out.shape[0]
    for t in range(trials):
        tok, _ = beam_decode(W, dims, spec, enc_out, max_new_tokens, beam_size=beam_size, logit_noise=noise,
                             noise_seed=1000 + t)
        for b in range(enc_out.shape[0]):
            stable[b] = stable[b] and bool(torch.equal(tok[b], ref_tokens[b]))
    return stable
