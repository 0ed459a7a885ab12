import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
from janus_b200.whisper.config import PRESETS, tokenizer_spec
from janus_b200.whisper.engine import DecodeSession, WhisperEngine
from janus_b200.whisper.weights import init_random_weights
from oracle import whisper_ref as R
for preset in ("base.en", "tiny", "small"):
    dims = PRESETS[preset]; spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=21, profile="confident")
    eng = WhisperEngine(dims, W, torch.device("cuda", 0))
    B, n_new = 3, 10
    g = torch.Generator().manual_seed(13)
    enc32 = torch.nn.functional.layer_norm(torch.randn(B, dims.n_audio_ctx, dims.n_audio_state, generator=g), (dims.n_audio_state,)).bfloat16().float()
    ref_tok, ref_lp, margins = R.greedy_decode(W, dims, spec, enc32, n_new, return_margins=True)
    st = R.DecoderState(W, dims, enc32)
    full = torch.cat([torch.tensor([list(spec.sot_sequence)] * B), ref_tok[:, :-1]], 1)
    ref_logits = st.step(full)[:, -1]
    enc = enc32.bfloat16().cuda()
    for latent in (True, False):
        ses = DecodeSession(eng, batch=B, max_new_tokens=n_new, latent=latent, use_graph=False)
        ses.decode(enc, forced=ref_tok.cuda()); torch.cuda.synchronize()
        err = (ses.logits[:, :dims.n_vocab].cpu() - ref_logits).abs()
        ch = ses.choice[:, ses.n_prompt:].cpu().long()
        print(f"{preset} latent={latent}: logit err max {err.max():.4f} mean {err.mean():.5f}; argmax agree {(ch == ref_tok).float().mean():.3f}; min margin where differ {margins[ch != ref_tok].min() if (ch != ref_tok).any() else -1}")
    # kernel-level check at this d, T=1500
    d, H, T, n = dims.n_audio_state, dims.n_text_head, 1500, 3
    qlat = torch.zeros(n, 16, d, device="cuda", dtype=torch.bfloat16)
    qlat[:, :H] = (torch.randn(n, H, d, device="cuda") * (8.0 / d ** 0.5)).bfloat16()
    z = torch.full((n, H, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    _abi.call("janus_dec_xlat_bf16", qlat.data_ptr(), enc.data_ptr(), z.data_ptr(), n, T, d, H, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    X = enc.float()
    s = torch.einsum("ntd,nhd->nht", X, qlat[:, :H].float()) * 0.125
    ref = torch.einsum("nht,ntd->nhd", torch.softmax(s, -1), X)
    print(f"   xlat kernel d={d} H={H} T={T}: max rel err {((z.float() - ref).abs() / ref.abs().amax(-1, keepdim=True)).max():.4f}")
