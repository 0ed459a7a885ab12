"""Latent vs K/V decode sessions against the fp32 oracle at Whisper-small shapes (diagnostic)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200.whisper.config import PRESETS, tokenizer_spec
from janus_b200.whisper.engine import DecodeSession, WhisperEngine
from janus_b200.whisper.weights import init_random_weights
from oracle import whisper_ref as R
dims = PRESETS["small"]; spec = tokenizer_spec(dims)
B, n_new = 4, 5
for profile in ("diverse", "confident"):
    W = init_random_weights(dims, seed=5, profile=profile)
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    for kind in ("randn", "ln"):
        g = torch.Generator().manual_seed(11)
        enc32 = torch.randn(B, dims.n_audio_ctx, dims.n_audio_state, generator=g)
        if kind == "ln":
            enc32 = torch.nn.functional.layer_norm(enc32 * torch.linspace(0.2, 3, dims.n_audio_state), (dims.n_audio_state,))
        enc32 = enc32.bfloat16().float()
        ref_tok, ref_lp, margins = R.greedy_decode(W, dims, spec, enc32, n_new, return_margins=True)
        st = R.DecoderState(W, dims, enc32)
        full = torch.cat([torch.tensor([list(spec.sot_sequence)] * B), ref_tok[:, :-1]], 1)
        ref_logits = st.step(full)[:, -1]
        enc = enc32.bfloat16().cuda()
        for latent in (False, True):
            ses = DecodeSession(eng, batch=B, max_new_tokens=n_new, latent=latent, use_graph=False)
            ses.decode(enc, forced=ref_tok.cuda()); torch.cuda.synchronize()
            got = ses.logits[:, :dims.n_vocab].cpu()
            err = (got - ref_logits).abs()
            ch = ses.choice[:, ses.n_prompt:].cpu().long()
            print(f"{profile:9s} enc={kind:5s} latent={latent}: logit err max {err.max():.4f} mean {err.mean():.5f} "
                  f"(std {ref_logits.std():.3f}) argmax agree {(ch == ref_tok).float().mean():.3f}")
