"""Diagnoses free-running divergences of bench windows against the oracle: for each requested window prints, at the
first divergence, the oracle's top-1 / top-2 margin AFTER the rules and the gap of the rule decision itself
(|logsumexp(timestamps) - max(text log-prob)|: 'timestamps outweigh text' -> sample a timestamp)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from janus_b200.pipeline import SemanticEncoder
from janus_b200.synth import make_clips
from janus_b200.whisper.weights import init_random_weights
from oracle import whisper_ref as R
from oracle.logmel_ref import log_mel_window

WIN = [int(v) for v in os.environ.get("WIN", "10,13,0,4").split(",")]
B, N = 128, 480000
enc = SemanticEncoder("small", batch=B, max_new_tokens=128, lane_batch=B, n_lanes=1)
x = make_clips(B, 30.0, 16000, seed=1234, device="cuda", chunk=64).reshape(-1)
res = enc.encode_device(x, (N,) * B).synchronize()
got = res.tokens.cpu().long()
dims, spec = enc.dims, enc.spec
W = init_random_weights(dims, seed=0, profile="confident")
sup = R.suppress_mask_for(spec)
tb = spec.timestamp_begin
for w in WIN:
    clip = x[w * N:(w + 1) * N].cpu().numpy()
    with torch.no_grad():
        mel = torch.from_numpy(log_mel_window(clip, dims.n_mels))[None]
        e = R.encode(W, dims, mel)
        tok_s, _, m_s = R.greedy_decode(W, dims, spec, e, 128, return_margins=True)
        tok_v, _, m_v = R.greedy_decode(W, dims, spec, e, 128, return_margins=True, vectorized=True)
    print(f"window {w}: scalar == vectorized oracle: {bool(torch.equal(tok_s, tok_v))}")
    neq = (got[w] != tok_s[0]).nonzero()
    if neq.numel() == 0:
        print(f"  bit-exact; min margin {float(m_s[0].min()):.4f}")
        continue
    t = int(neq[0])
    # recompute the oracle's logits at step t
    st = R.DecoderState(W, dims, e)
    full = torch.tensor([list(spec.sot_sequence) + tok_s[0, :t].tolist()])
    with torch.no_grad():
        lg = st.step(full)[0, -1]
    hist = tok_s[0, :t].tolist()
    # the rule decision: filtered logits BEFORE the ts-vs-text rule
    f = R.apply_logit_rules(lg, hist, spec, sup)
    _, gap = R.apply_logit_rules(lg, hist, spec, sup, return_gap=True)
    print(f"  gap of the timestamps-vs-text rule decision at that step: {gap:.4f}")
    print(f"  first divergence at token {t}: oracle {int(tok_s[0, t])} (margin after rules {float(m_s[0, t]):.4f}), CUDA {int(got[w, t])};"
          f" oracle token is {'timestamp' if int(tok_s[0, t]) >= tb else 'text'}, CUDA token is {'timestamp' if int(got[w, t]) >= tb else 'text'}")
    print(f"  oracle raw logit of its token {float(lg[int(tok_s[0, t])]):.4f}, of CUDA's token {float(lg[int(got[w, t])]):.4f}; "
          f"forbidden for the oracle: {bool(torch.isinf(f[int(got[w, t])]))}")
