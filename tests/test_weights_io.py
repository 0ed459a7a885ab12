"""Checkpoint loaders (janus_b200/whisper/weights.py: load_weights and the three key maps) and the tokenizer hook
(janus_b200/whisper/tokenizer.py) -- CPU only. What the reference does here: `WhisperModel(model_size, ...)` loads a
pretrained checkpoint (/root/reference/backend/services/transcriber.py:23-27) and `segment.text` comes out of its
tokenizer (:59-64). No checkpoint is downloadable in this image, so the loaders are exercised on a randomly
initialised Hugging Face `WhisperForConditionalGeneration` -- an INDEPENDENT implementation: its fp32 forward on the
loaded weights must equal the oracle's forward on the converted dict."""
import os

import numpy as np
import pytest
import torch

from janus_b200.whisper import weights as Wt
from janus_b200.whisper.config import PRESETS, mini_dims, tokenizer_spec
from oracle import whisper_ref as R

transformers = pytest.importorskip("transformers")


def _hf_model(d=64, heads=1, layers=2, n_mels=80, vocab=51865, seed=0):
    torch.manual_seed(seed)
    cfg = transformers.WhisperConfig(
        vocab_size=vocab, num_mel_bins=n_mels, d_model=d, encoder_layers=layers, decoder_layers=layers,
        encoder_attention_heads=heads, decoder_attention_heads=heads, encoder_ffn_dim=4 * d, decoder_ffn_dim=4 * d,
        max_source_positions=1500, max_target_positions=448, activation_function="gelu", attn_implementation="eager")
    m = transformers.WhisperForConditionalGeneration(cfg).eval()
    with torch.no_grad():   # HF initialises biases / LayerNorms to constants: randomise them so a swap would show
        for n, p in m.named_parameters():
            if p.dim() == 1:
                p.add_(0.1 * torch.randn_like(p))
    return m


def _check_against_hf(m, dims, W):
    spec = tokenizer_spec(dims)
    mel = torch.randn(2, dims.n_mels, 3000, generator=torch.Generator().manual_seed(1))
    toks = torch.tensor([[spec.sot, 50259, 50359, 50364, 1000, 2000], [spec.sot, 50259, 50359, 50365, 7, 9]])
    with torch.no_grad():
        enc_hf = m.model.encoder(mel).last_hidden_state
        logits_hf = m(input_features=mel, decoder_input_ids=toks).logits
        enc = R.encode(W, dims, mel)
        logits = R.DecoderState(W, dims, enc).step(toks)
    assert (enc - enc_hf).abs().max() < 2e-4
    assert (logits - logits_hf).abs().max() < 2e-3


def test_hf_state_dict_and_safetensors_dir_roundtrip(tmp_path):
    m = _hf_model()
    W = Wt.from_hf_state_dict(m.state_dict())
    dims = Wt.dims_from_weights(W, n_heads=1)
    assert (dims.n_audio_state, dims.n_audio_layer, dims.n_vocab, dims.n_mels) == (64, 2, 51865, 80)
    _check_against_hf(m, dims, W)
    m.save_pretrained(str(tmp_path / "hf"), safe_serialization=True)
    dims2, W2 = Wt.load_weights(str(tmp_path / "hf"))
    assert dims2.n_audio_head == 1 and dims2.n_text_layer == 2     # head count from config.json
    assert set(W2) == set(W) and all(torch.equal(W[k], W2[k]) for k in W)


def test_openai_checkpoint_roundtrip(tmp_path):
    """OpenAI `<size>.pt` layout ({'dims': ..., 'model_state_dict': ...}, key names of whisper/model.py). The key names
    are restated from the published module structure (encoder.blocks.N.attn.query ..., mlp.0 / mlp.2)."""
    dims = mini_dims(d=64, heads=1, layers=2)
    W = Wt.init_random_weights(dims, seed=5)
    sd = {}
    for c in ("conv1", "conv2"):
        sd[f"encoder.{c}.weight"], sd[f"encoder.{c}.bias"] = W[f"enc.{c}.w"], W[f"enc.{c}.b"]
    sd["encoder.positional_embedding"] = W["enc.pos"]
    sd["decoder.token_embedding.weight"], sd["decoder.positional_embedding"] = W["dec.tok_emb"], W["dec.pos"]
    for side, pre, n in (("encoder", "enc", dims.n_audio_layer), ("decoder", "dec", dims.n_text_layer)):
        for i in range(n):
            o, p = f"{side}.blocks.{i}", f"{pre}.l{i}"
            blocks = [("attn", "attn", "attn_ln")] + ([("cross_attn", "cross", "cross_attn_ln")] if side == "decoder" else [])
            for oa, pa, oln in blocks:
                for a, b in (("query", "q"), ("key", "k"), ("value", "v"), ("out", "o")):
                    sd[f"{o}.{oa}.{a}.weight"] = W[f"{p}.{pa}.{b}.w"]
                    if b != "k":
                        sd[f"{o}.{oa}.{a}.bias"] = W[f"{p}.{pa}.{b}.b"]
                ln = "attn_ln" if pa == "attn" else "cross_ln"
                sd[f"{o}.{oln}.weight"], sd[f"{o}.{oln}.bias"] = W[f"{p}.{ln}.w"], W[f"{p}.{ln}.b"]
            sd[f"{o}.mlp_ln.weight"], sd[f"{o}.mlp_ln.bias"] = W[f"{p}.mlp_ln.w"], W[f"{p}.mlp_ln.b"]
            sd[f"{o}.mlp.0.weight"], sd[f"{o}.mlp.0.bias"] = W[f"{p}.mlp.fc1.w"], W[f"{p}.mlp.fc1.b"]
            sd[f"{o}.mlp.2.weight"], sd[f"{o}.mlp.2.bias"] = W[f"{p}.mlp.fc2.w"], W[f"{p}.mlp.fc2.b"]
    sd["encoder.ln_post.weight"], sd["encoder.ln_post.bias"] = W["enc.ln_post.w"], W["enc.ln_post.b"]
    sd["decoder.ln.weight"], sd["decoder.ln.bias"] = W["dec.ln.w"], W["dec.ln.b"]
    path = str(tmp_path / "mini.pt")
    torch.save({"dims": {"n_audio_head": 1, "n_text_head": 1}, "model_state_dict": {k: v.half() for k, v in sd.items()}}, path)
    dims2, W2 = Wt.load_weights(path)
    assert dims2.n_audio_head == 1 and dims2.n_audio_layer == 2
    assert set(W2) == set(W)
    for k in W:
        assert torch.equal(W2[k], W[k].half().float()), k


def test_ct2_model_bin_roundtrip(tmp_path):
    """CTranslate2 `model.bin` (what faster-whisper loads): fused q/k/v and k/v projections, an int8 weight with its
    per-row scale, aliases. Reader checked against the writer only (the format is recalled; see weights.py)."""
    dims = mini_dims(d=64, heads=1, layers=1)
    W = Wt.init_random_weights(dims, seed=6)
    V = {}
    for c in ("conv1", "conv2"):
        V[f"encoder/{c}/weight"], V[f"encoder/{c}/bias"] = W[f"enc.{c}.w"].half(), W[f"enc.{c}.b"].half()
    V["encoder/position_encodings/encodings"] = W["enc.pos"]
    V["decoder/embeddings/weight"] = W["dec.tok_emb"].half()
    V["decoder/position_encodings/encodings"] = W["dec.pos"].half()
    z = torch.zeros(64)

    def ln(cp, p):
        V[cp + "/gamma"], V[cp + "/beta"] = W[p + ".w"].half(), W[p + ".b"].half()

    def lin(cp, w, b):
        V[cp + "/weight"], V[cp + "/bias"] = w.half(), b.half()

    for side, pre in (("encoder", "enc"), ("decoder", "dec")):
        cp, p = f"{side}/layer_0", f"{pre}.l0"
        ln(cp + "/self_attention/layer_norm", p + ".attn_ln")
        lin(cp + "/self_attention/linear_0", torch.cat([W[p + ".attn.q.w"], W[p + ".attn.k.w"], W[p + ".attn.v.w"]]),
            torch.cat([W[p + ".attn.q.b"], z, W[p + ".attn.v.b"]]))
        lin(cp + "/self_attention/linear_1", W[p + ".attn.o.w"], W[p + ".attn.o.b"])
        if side == "decoder":
            ln(cp + "/attention/layer_norm", p + ".cross_ln")
            lin(cp + "/attention/linear_0", W[p + ".cross.q.w"], W[p + ".cross.q.b"])
            lin(cp + "/attention/linear_1", torch.cat([W[p + ".cross.k.w"], W[p + ".cross.v.w"]]),
                torch.cat([z, W[p + ".cross.v.b"]]))
            lin(cp + "/attention/linear_2", W[p + ".cross.o.w"], W[p + ".cross.o.b"])
        ln(cp + "/ffn/layer_norm", p + ".mlp_ln")
        lin(cp + "/ffn/linear_0", W[p + ".mlp.fc1.w"], W[p + ".mlp.fc1.b"])
        lin(cp + "/ffn/linear_1", W[p + ".mlp.fc2.w"], W[p + ".mlp.fc2.b"])
    ln("encoder/layer_norm", "enc.ln_post")
    ln("decoder/layer_norm", "dec.ln")
    # one int8-quantised weight: q = round(w * scale), scale = 127 / max|row|
    w = W["enc.l0.mlp.fc1.w"]
    scale = 127.0 / w.abs().amax(dim=1)
    V["encoder/layer_0/ffn/linear_0/weight"] = torch.round(w * scale[:, None]).to(torch.int8)
    V["encoder/layer_0/ffn/linear_0/weight_scale"] = scale
    d = tmp_path / "ct2"
    d.mkdir()
    Wt.write_ct2_variables(str(d / "model.bin"), V, aliases=[("decoder/projection/weight", "decoder/embeddings/weight")])
    raw = Wt.read_ct2_variables(str(d / "model.bin"))
    assert torch.equal(raw["decoder/projection/weight"], raw["decoder/embeddings/weight"])
    dims2, W2 = Wt.load_weights(str(d))
    assert (dims2.n_audio_state, dims2.n_audio_layer, dims2.n_text_layer) == (64, 1, 1)
    assert set(W2) == set(W)
    for k in W:
        tol = 2e-2 if k == "enc.l0.mlp.fc1.w" else 2e-3
        assert (W2[k] - W[k]).abs().max() <= tol * max(1.0, float(W[k].abs().max())), k
    assert "enc.l0.attn.k.b" not in W2 and torch.equal(W2["dec.l0.cross.v.b"], W["dec.l0.cross.v.b"].half().float())


def _synthetic_tokenizer(path):
    """A byte-level BPE vocabulary with Whisper's special-token NAMES appended (ids differ from the real vocabulary:
    the hook must look them up by name)."""
    from tokenizers import AddedToken, Tokenizer, decoders, models, pre_tokenizers, trainers
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    tok.decoder = decoders.ByteLevel()
    tr = trainers.BpeTrainer(vocab_size=400, special_tokens=[], initial_alphabet=pre_tokenizers.ByteLevel.alphabet())
    tok.train_from_iterator(["hello world this is a test of the janus packet encoder (music) [noise] -- hi"] * 10, tr)
    specials = ["<|endoftext|>", "<|startoftranscript|>", "<|en|>", "<|translate|>", "<|transcribe|>", "<|startoflm|>",
                "<|startofprev|>", "<|nospeech|>", "<|notimestamps|>"] + [f"<|{i * 0.02:.2f}|>" for i in range(1501)]
    tok.add_special_tokens([AddedToken(s, special=True) for s in specials])
    tok.save(path)
    return tok


def test_tokenizer_hook(tmp_path):
    from janus_b200.pipeline import join_segments, split_text_segments
    from janus_b200.whisper.tokenizer import WhisperTokenizer
    path = str(tmp_path / "tokenizer.json")
    tok = _synthetic_tokenizer(path)
    wt = WhisperTokenizer.from_file(path)
    spec = wt.spec()
    n = tok.token_to_id("<|endoftext|>")
    assert (spec.eot, spec.sot, spec.lang_en, spec.transcribe) == (n, n + 1, n + 2, n + 4)
    assert spec.timestamp_begin == tok.token_to_id("<|0.00|>") and spec.no_speech == n + 7
    assert spec.sot_sequence == (n + 1, n + 2, n + 4)
    assert spec.blank == tok.encode(" ").ids[0]
    # byte-level alphabet: the single-character symbols get the same ids as in GPT-2 / Whisper's vocabulary, so the
    # computed non-speech set starts like the published list in config.py
    from janus_b200.whisper.config import _NON_SPEECH_MULTI
    assert set(_NON_SPEECH_MULTI[:23]) <= set(spec.non_speech)
    ids = tok.encode(" hello world").ids
    ts = spec.timestamp_begin
    row = [ts, *ids, ts + 50, ts + 50, *tok.encode(" hi").ids, ts + 80, spec.eot, spec.eot]
    assert join_segments(split_text_segments(row, spec)) == "hello world hi"
    assert all(s in spec.suppress_ids for s in (spec.sot, spec.no_speech, spec.no_timestamps))
