"""Pins the STT oracles (no GPU): the numpy log-mel restatement against the independent HF
WhisperFeatureExtractor, and the fp32 Whisper restatement against HF transformers' Whisper with the
same weights. (faster-whisper / CTranslate2 themselves are not installable offline: parity unpinned.)"""
import numpy as np
import pytest
import torch

from janus_b200.synth import make_clips
from janus_b200.whisper.config import PRESETS, detokenize, mini_dims, tokenizer_spec
from janus_b200.whisper.weights import init_random_weights
from oracle import whisper_ref as R
from oracle.logmel_ref import log_mel_utterance, log_mel_window, mel_filters

transformers = pytest.importorskip("transformers")


def test_logmel_matches_hf_feature_extractor():
    x = make_clips(1, 30.0, 16000, seed=3).numpy()[0]
    ours = log_mel_window(x, 80)
    fe = transformers.WhisperFeatureExtractor(feature_size=80)
    hf = fe(x, sampling_rate=16000, return_tensors="np").input_features[0]
    assert ours.shape == hf.shape == (80, 3000)
    # identical away from the tail (HF reflects at 30 s, faster-whisper sees 160 zeros first)
    assert np.abs(ours[:, :2990] - hf[:, :2990]).max() < 5e-5
    assert np.abs(fe.mel_filters.T - mel_filters(80)).max() < 1e-6


def test_logmel_window_semantics():
    x = make_clips(1, 30.0, 16000, seed=3).numpy()[0]
    full = log_mel_utterance(x[:80000], 80)
    assert full.shape == (80, 501)                      # (80000 + 160) // 160 frames kept
    win = log_mel_window(x[:80000], 80)
    assert win.shape == (80, 3000)
    assert np.array_equal(win[:, :500], full[:, :500])  # content_frames = 500
    assert (win[:, 500:] == 0).all()                    # zero padding in FEATURE space
    assert full.min() >= full.max() - 2.0 - 1e-6        # (max - 8 + 4)/4 clamp -> dynamic range 2.0


def _to_hf(W, dims):
    cfg = transformers.WhisperConfig(
        vocab_size=dims.n_vocab, num_mel_bins=dims.n_mels, d_model=dims.d, encoder_layers=dims.n_audio_layer,
        decoder_layers=dims.n_text_layer, encoder_attention_heads=dims.n_audio_head,
        decoder_attention_heads=dims.n_text_head, encoder_ffn_dim=4 * dims.d, decoder_ffn_dim=4 * dims.d,
        max_source_positions=dims.n_audio_ctx, max_target_positions=dims.n_text_ctx,
        activation_function="gelu", attn_implementation="eager")
    m = transformers.WhisperForConditionalGeneration(cfg).eval()
    sd = {}

    def ln(hp, p):
        sd[hp + ".weight"], sd[hp + ".bias"] = W[p + ".w"], W[p + ".b"]

    def attn(hp, p):
        for a, b in (("q_proj", "q"), ("k_proj", "k"), ("v_proj", "v"), ("out_proj", "o")):
            sd[f"{hp}.{a}.weight"] = W[f"{p}.{b}.w"]
            if b != "k":
                sd[f"{hp}.{a}.bias"] = W[f"{p}.{b}.b"]

    def mlp(hp, p):
        sd[hp + ".fc1.weight"], sd[hp + ".fc1.bias"] = W[p + ".mlp.fc1.w"], W[p + ".mlp.fc1.b"]
        sd[hp + ".fc2.weight"], sd[hp + ".fc2.bias"] = W[p + ".mlp.fc2.w"], W[p + ".mlp.fc2.b"]

    sd["model.encoder.conv1.weight"], sd["model.encoder.conv1.bias"] = W["enc.conv1.w"], W["enc.conv1.b"]
    sd["model.encoder.conv2.weight"], sd["model.encoder.conv2.bias"] = W["enc.conv2.w"], W["enc.conv2.b"]
    sd["model.encoder.embed_positions.weight"] = W["enc.pos"]
    for i in range(dims.n_audio_layer):
        hp, p = f"model.encoder.layers.{i}", f"enc.l{i}"
        attn(hp + ".self_attn", p + ".attn"); ln(hp + ".self_attn_layer_norm", p + ".attn_ln")
        mlp(hp, p); ln(hp + ".final_layer_norm", p + ".mlp_ln")
    ln("model.encoder.layer_norm", "enc.ln_post")
    sd["model.decoder.embed_tokens.weight"] = W["dec.tok_emb"]
    sd["model.decoder.embed_positions.weight"] = W["dec.pos"]
    for i in range(dims.n_text_layer):
        hp, p = f"model.decoder.layers.{i}", f"dec.l{i}"
        attn(hp + ".self_attn", p + ".attn"); ln(hp + ".self_attn_layer_norm", p + ".attn_ln")
        attn(hp + ".encoder_attn", p + ".cross"); ln(hp + ".encoder_attn_layer_norm", p + ".cross_ln")
        mlp(hp, p); ln(hp + ".final_layer_norm", p + ".mlp_ln")
    ln("model.decoder.layer_norm", "dec.ln")
    sd["proj_out.weight"] = W["dec.tok_emb"]
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not missing and not unexpected
    return m


def test_whisper_oracle_matches_hf_transformers():
    dims = mini_dims(d=128, heads=2, layers=2)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=1, profile="diverse")
    torch.manual_seed(0)
    mel = torch.randn(2, 80, 3000) * 0.5
    enc = R.encode(W, dims, mel)
    hf = _to_hf(W, dims)
    with torch.no_grad():
        henc = hf.model.encoder(mel).last_hidden_state
        assert float((henc - enc).abs().max()) < 2e-5
        toks = torch.tensor([[spec.sot, spec.lang_en, spec.transcribe, spec.timestamp_begin, 100, 2000, 31000]] * 2)
        hl = hf(input_features=mel, decoder_input_ids=toks).logits
    st = R.DecoderState(W, dims, enc)
    ol = torch.cat([st.step(toks[:, :3]), st.step(toks[:, 3:4]), st.step(toks[:, 4:])], 1)  # KV-cache path
    assert float((hl - ol).abs().max()) < 2e-3 * max(1.0, float(ol.abs().max()))


def test_greedy_rules():
    dims = mini_dims(d=128, heads=2, layers=2)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=3, profile="diverse")
    mel = torch.randn(1, 80, 3000) * 0.5
    out, lp, mg = R.greedy_decode(W, dims, spec, R.encode(W, dims, mel), 24, return_margins=True)
    seq = out[0].tolist()
    tb = spec.timestamp_begin
    assert tb <= seq[0] <= tb + spec.max_initial_timestamp_index      # first token: timestamp <= 1.0 s
    assert seq[1] < tb                                                 # ... followed by text (or eot)
    stamps = [t for t in seq if t >= tb]
    assert stamps == sorted(stamps)                                    # timestamps never decrease
    sup = set(spec.suppress_ids)
    assert not any(t in sup for t in seq)
    assert detokenize(seq, spec).count(" w") == sum(1 for t in seq if t < spec.eot)


def test_presets_param_counts():
    # architecture sanity: parameter counts of the named models (tiny 37.8 M, small ~242 M)
    for name, lo, hi in (("tiny", 37e6, 39e6), ("small", 240e6, 245e6)):
        d = PRESETS[name]
        n = d.n_audio_state
        enc = d.n_mels * n * 3 + n + 3 * n * n + n + d.n_audio_layer * (4 * n * n + 3 * n + 8 * n * n + 5 * n + 4 * n) + 2 * n
        dec = d.n_vocab * n + 448 * n + d.n_text_layer * (8 * n * n + 6 * n + 8 * n * n + 5 * n + 6 * n) + 2 * n
        assert lo < enc + dec < hi, (name, enc + dec)


def test_beam_oracle_beam1_is_greedy_and_beam5_never_scores_lower():
    """oracle/whisper_ref.py beam_decode: beam 1 reproduces greedy_decode; the best of 5 beams has a normalised
    score >= the greedy hypothesis (same length here: random weights do not emit eot)."""
    import torch
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    dims = mini_dims(d=64, heads=1, layers=1)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=5, profile="diverse")
    torch.manual_seed(0)
    enc = torch.randn(2, dims.n_audio_ctx, dims.n_audio_state) * 0.5
    with torch.no_grad():
        g, glp = R.greedy_decode(W, dims, spec, enc, 6)
        b1, b1lp = R.beam_decode(W, dims, spec, enc, 6, beam_size=1)
        b5, b5lp = R.beam_decode(W, dims, spec, enc, 6, beam_size=5)
    assert torch.equal(g, b1) and torch.allclose(glp, b1lp, atol=1e-4)
    assert bool((b5lp >= glp - 1e-4).all())


def test_beam_step_select_walk():
    """The published walk: eot candidates finish, the first `beam` others continue, -inf is no candidate."""
    from oracle.whisper_ref import beam_step_select
    eot = 9
    ninf = float("-inf")
    cand = [(-1.0, 0, 3), (-1.5, 0, eot), (-2.0, 0, 4), (-0.5, 1, eot), (-1.2, 1, 7), (-3.0, 1, 2),
            (ninf, 2, 1), (ninf, 2, eot), (ninf, 2, 5)]
    nxt, fin = beam_step_select(cand, 3, eot)
    assert fin == [(-0.5, 1), (-1.5, 0)]
    assert nxt == [(-1.0, 0, 3), (-1.2, 1, 7), (-2.0, 0, 4)]
    # fewer live candidates than beams: the remaining slots are dead beams
    nxt, fin = beam_step_select([(-1.0, 0, 3), (ninf, 1, 4), (ninf, 1, 5)], 2, eot)
    assert nxt == [(-1.0, 0, 3), (ninf, 1, eot)] and fin == []


def test_decode_oracle_reproduces_golden_vectors():
    """tests/golden/decode_golden.json (greedy + beam-5 decodes, generated by tests/golden/make_golden.py) pins the
    Whisper oracle against drift: same seeds -> same tokens, finished lists and scores."""
    import json
    import os
    import sys
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, here)
    try:
        import make_golden
    finally:
        sys.path.remove(here)
    gold = json.load(open(os.path.join(here, "decode_golden.json")))
    now = make_golden.decode_cases(gold["n_new"], gold["beam"])
    for name, g in gold["cases"].items():
        n = now["cases"][name]
        assert n["greedy_tokens"] == g["greedy_tokens"], name
        assert n["beam_tokens"] == g["beam_tokens"], name
        assert n["beam_decisive"] == g["beam_decisive"], name
        for a, b in zip(n["beam_sum_logprob"] + n["greedy_sum_logprob"], g["beam_sum_logprob"] + g["greedy_sum_logprob"]):
            assert abs(a - b) <= 2e-3
        for fa, fb in zip(n["finished"], g["finished"]):
            assert [f["tokens"] for f in fa] == [f["tokens"] for f in fb] and [f["len"] for f in fa] == [f["len"] for f in fb]


def test_beam_oracle_matches_hf_transformers_beam_search():
    """Independent cross-check of the oracle's beam search (beam pruning over beam x vocab, cumulative log-probs of the
    RENORMALISED rule-filtered distribution, ranking by sum / length): HF transformers' generic `generate(num_beams=5)`
    on the same weights, with the oracle's logit rules plugged in as a LogitsProcessor that renormalises (HF applies
    processors after its own log-softmax and would otherwise accumulate unnormalised scores, unlike OpenAI Whisper /
    CTranslate2). Random weights do not emit <|endoftext|>, so this pins the pruning / ranking, not the
    finished-hypothesis bookkeeping (which differs between HF and OpenAI by design)."""
    import torch
    from transformers.generation import GenerationMixin, LogitsProcessor, LogitsProcessorList
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    dims = mini_dims(d=64, heads=1, layers=1)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=5, profile="diverse")
    hf = _to_hf(W, dims)
    torch.manual_seed(0)
    mel = torch.randn(2, dims.n_mels, 3000) * 0.5
    n_new, K = 8, 5
    with torch.no_grad():
        tok, lp = R.beam_decode(W, dims, spec, R.encode(W, dims, mel), n_new, beam_size=K)
    sup = R.suppress_mask_for(spec)
    n_prompt = len(spec.sot_sequence)

    class Rules(LogitsProcessor):
        def __call__(self, input_ids, scores):
            out = scores.clone()
            for r in range(scores.shape[0]):
                out[r] = torch.log_softmax(R.apply_logit_rules(scores[r], input_ids[r, n_prompt:].tolist(), spec, sup), -1)
            return out

    gc = transformers.GenerationConfig(num_beams=K, max_new_tokens=n_new, do_sample=False, length_penalty=1.0,
                                       early_stopping=False, eos_token_id=spec.eot, pad_token_id=spec.eot,
                                       decoder_start_token_id=spec.sot_sequence[0], num_return_sequences=1,
                                       output_scores=True, return_dict_in_generate=True)
    prompt = torch.tensor([list(spec.sot_sequence)] * mel.shape[0])
    with torch.no_grad():
        out = GenerationMixin.generate(hf, input_features=mel, decoder_input_ids=prompt, generation_config=gc,
                                       attention_mask=None, logits_processor=LogitsProcessorList([Rules()]))
    assert out.sequences[:, n_prompt:].tolist() == tok.tolist()
    assert torch.allclose(out.sequences_scores, lp / n_new, atol=1e-3)


@pytest.mark.parametrize("vocab", [51865, 51864])
def test_logit_rules_match_hf_whisper_logits_processors(vocab):
    """The decode-time rules (SuppressBlank, SuppressTokens, ApplyTimestampRules incl. max_initial_timestamp and the
    "timestamps outweigh every text token" rule) are recalled semantics of faster-whisper / openai-whisper.
    HF transformers ships an independent implementation of the same rules (SuppressTokensAtBegin + SuppressTokens +
    WhisperTimeStampLogitsProcessor): on arbitrary token histories and logits the two must forbid exactly the same
    tokens. The CUDA kernels (dec_pick_kernel, dec_beam_cand_kernel) are tested against the oracle's rules."""
    import numpy as np
    import torch
    from transformers.generation.logits_process import (SuppressTokensAtBeginLogitsProcessor, SuppressTokensLogitsProcessor,
                                                        WhisperTimeStampLogitsProcessor)
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from oracle import whisper_ref as R
    spec = tokenizer_spec(mini_dims(vocab=vocab))
    n_prompt = len(spec.sot_sequence)
    gc = transformers.GenerationConfig(eos_token_id=spec.eot, no_timestamps_token_id=spec.no_timestamps,
                                       max_initial_timestamp_index=spec.max_initial_timestamp_index)
    chain = [SuppressTokensAtBeginLogitsProcessor([spec.blank, spec.eot], n_prompt),
             SuppressTokensLogitsProcessor(list(spec.suppress_ids)),
             WhisperTimeStampLogitsProcessor(gc, n_prompt)]
    sup = R.suppress_mask_for(spec)
    rng = np.random.default_rng(0)
    tb = spec.timestamp_begin
    fired = 0
    for _ in range(100):
        gen = []
        for _ in range(int(rng.integers(0, 9))):   # arbitrary mixtures of text tokens, timestamps and specials
            r = rng.random()
            if r < 0.45:
                gen.append(int(rng.integers(0, spec.eot)))
            elif r < 0.9:
                gen.append(int(rng.integers(tb, spec.n_vocab)))
            else:
                gen.append(int(rng.integers(spec.eot, tb)))
        logits = torch.from_numpy(rng.standard_normal(spec.n_vocab).astype(np.float32)) * float(rng.choice([0.5, 3.0, 10.0]))
        if rng.random() < 0.3:
            logits[tb:] += 4.0   # lets the "timestamps outweigh text" rule fire
        ours = R.apply_logit_rules(logits, gen, spec, sup)
        sc = logits.clone()[None]
        ids = torch.tensor([list(spec.sot_sequence) + gen])
        for p in chain:
            sc = p(ids, sc)
        assert torch.equal(torch.isinf(ours), torch.isinf(sc[0])), gen
        fired += int(bool(torch.isinf(ours[:tb]).all()) and len(gen) > 0)
    assert fired > 0


def test_greedy_oracle_matches_hf_generate_end_to_end():
    """Whole greedy decode, oracle vs HF: HF's Whisper model (same weights) + HF's own Whisper logits processors +
    HF's greedy loop must emit the oracle's tokens (fp32 both; asserted where the oracle's top-1 margin > 1e-3)."""
    import torch
    from transformers.generation import GenerationMixin, LogitsProcessorList
    from transformers.generation.logits_process import (SuppressTokensAtBeginLogitsProcessor, SuppressTokensLogitsProcessor,
                                                        WhisperTimeStampLogitsProcessor)
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    dims = mini_dims(d=64, heads=1, layers=2)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=9, profile="diverse")
    hf = _to_hf(W, dims)
    torch.manual_seed(1)
    mel = torch.randn(3, dims.n_mels, 3000) * 0.5
    n_new, n_prompt = 20, len(spec.sot_sequence)
    with torch.no_grad():
        tok, lp, margins = R.greedy_decode(W, dims, spec, R.encode(W, dims, mel), n_new, return_margins=True)
    gc = transformers.GenerationConfig(eos_token_id=spec.eot, pad_token_id=spec.eot, no_timestamps_token_id=spec.no_timestamps,
                                       max_initial_timestamp_index=spec.max_initial_timestamp_index, num_beams=1,
                                       do_sample=False, max_new_tokens=n_new, decoder_start_token_id=spec.sot_sequence[0])
    chain = LogitsProcessorList([SuppressTokensAtBeginLogitsProcessor([spec.blank, spec.eot], n_prompt),
                                 SuppressTokensLogitsProcessor(list(spec.suppress_ids)),
                                 WhisperTimeStampLogitsProcessor(gc, n_prompt)])
    prompt = torch.tensor([list(spec.sot_sequence)] * mel.shape[0])
    with torch.no_grad():
        out = GenerationMixin.generate(hf, input_features=mel, decoder_input_ids=prompt, generation_config=gc,
                                       logits_processor=chain)
    got = out[:, n_prompt:]
    assert float(margins.min()) > 1e-3, "pick another seed: the oracle itself is at a tie"
    n = min(got.shape[1], n_new)
    assert got[:, :n].tolist() == tok[:, :n].tolist()


def test_vectorized_logit_rules_equal_the_scalar_rules():
    """apply_logit_rules_batch (used by the timed CPU baseline so that it is not handicapped by a per-sequence Python
    loop) forbids exactly the same tokens and keeps exactly the same values as apply_logit_rules, on random histories
    of both vocabulary layouts; greedy_decode(vectorized=True) returns the same tokens."""
    import random
    from janus_b200.whisper.config import PRESETS, tokenizer_spec
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    for name in ("tiny", "tiny.en"):
        dims = PRESETS[name]
        spec = tokenizer_spec(dims)
        sup = R.suppress_mask_for(spec)
        tb = spec.timestamp_begin
        rnd = random.Random(3)
        hs = []
        for _ in range(200):
            h = []
            for _ in range(rnd.choice([0, 1, 2, 3, 5, 9])):
                h.append(rnd.randint(tb, dims.n_vocab - 1) if rnd.random() < 0.4 else rnd.randint(0, spec.eot - 1))
            hs.append(h)
        logits = torch.randn(len(hs), dims.n_vocab, generator=torch.Generator().manual_seed(1)) * 3
        logits[::3, tb:] += 4          # rows where the timestamp mass outweighs every text token
        got = R.apply_logit_rules_batch(logits, hs, spec, sup)
        for b, h in enumerate(hs):
            ref = R.apply_logit_rules(logits[b], h, spec, sup)
            assert torch.equal(torch.isinf(ref), torch.isinf(got[b])), (name, b, h)
            keep = ~torch.isinf(ref)
            assert torch.equal(ref[keep], got[b][keep])
    dims = mini_dims(d=64, heads=1, layers=1)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=2, profile="confident")
    enc = torch.randn(3, dims.n_audio_ctx, dims.n_audio_state, generator=torch.Generator().manual_seed(4))
    with torch.no_grad():
        a, la = R.greedy_decode(W, dims, spec, enc, 20)
        b, lb = R.greedy_decode(W, dims, spec, enc, 20, vectorized=True)
    assert torch.equal(a, b) and torch.allclose(la, lb)
