"""Generates the committed golden fixtures. Run HERE (container with /root/reference mounted):

    python tests/golden/make_golden.py

* protocol_golden.json  -- produced by IMPORTING THE REFERENCE's own
  /root/reference/backend/common/protocol.py (msgpack is installed here), so these bytes pin the
  packet builder against the real reference.
* decode_golden.json    -- greedy and beam-5 decodes of the fp32 Whisper oracle (oracle/whisper_ref.py) on seeded
  weights / synthetic clips at a shrunken architecture: tokens, sums of log-probs, finished-hypothesis lists and
  whether each window's result is decisive (margins far above bf16 noise). faster-whisper / CTranslate2 are not
  installable here, so these vectors pin the CUDA decoder against the oracle ("parity unpinned").
* prosody_golden.npz    -- produced by the CPU oracle (oracle/yin_aubio.c). aubio itself is not
  installable here, so these vectors pin the CUDA kernels against the oracle, not against aubio
  ("parity unpinned" for per-frame f0).
"""
import enum
import importlib.util
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def load_reference_protocol():
    spec = importlib.util.spec_from_file_location(
        "ref_protocol", "/root/reference/backend/common/protocol.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def protocol_cases():
    class EmotionOverride(str, enum.Enum):  # shape of backend/api/types.py EmotionOverride
        AUTO = "auto"
    long_text = "the quick brown fox jumps over the lazy dog " * 8
    return [
        dict(text="Hello world", mode=0, prosody={'energy': 'Normal', 'pitch': 'High'}, override=None, ts=1234567890.0),
        dict(text="Hello world", mode=0, prosody={'energy': 'Normal', 'pitch': 'High'}, override="Relaxed", ts=1234567890.0),
        dict(text="hi", mode=0, prosody={'energy': 'Quiet', 'pitch': 'Deep'}, override=EmotionOverride.AUTO, ts=1.5),
        dict(text="x" * 40, mode=1, prosody={}, override="Auto", ts=0.0),
        dict(text="", mode=2, prosody={'energy': 'Loud', 'pitch': 'Normal'}, override=None, ts=1e9 + 0.125),
        dict(text=long_text.strip(), mode=0, prosody={'energy': 'Loud', 'pitch': 'Deep'}, override="Panicked", ts=1700000000.25),
        dict(text="café 你好 \U0001f600", mode=0, prosody={'energy': 'Normal', 'pitch': 'Normal'}, override=None, ts=42.0),
        dict(text="a" * 300, mode=0, prosody={'energy': 'Quiet', 'pitch': 'High'}, override=None, ts=3.0),
    ]


def make_protocol():
    ref = load_reference_protocol()
    out = []
    for c in protocol_cases():
        pkt = ref.JanusPacket(c["text"], ref.JanusMode(c["mode"]), c["prosody"], c["override"], c["ts"])
        ov = c["override"]
        out.append(dict(text=c["text"], mode=c["mode"], prosody=c["prosody"],
                        override=None if ov is None else str(ov.value if isinstance(ov, enum.Enum) else ov),
                        override_is_str_enum=isinstance(ov, enum.Enum), ts=c["ts"],
                        hex=pkt.serialize().hex()))
    with open(os.path.join(HERE, "protocol_golden.json"), "w") as f:
        json.dump(out, f, indent=1, ensure_ascii=True)
    print("protocol_golden.json", len(out), "cases")


def prosody_inputs():
    """name -> (sample_rate, list of consecutive utterances fed to ONE stateful extractor)."""
    from janus_b200.synth import make_clips, sine_wave
    rng = np.random.default_rng(7)
    clips16 = make_clips(3, 4.0, 16000, seed=1234).numpy()
    clips48 = make_clips(2, 2.0, 48000, seed=99).numpy()
    cases = {
        "ref_sine440_48k": (48000, [sine_wave(440.0, 0.5, 48000, 0.3)]),
        "ref_sine_quiet": (48000, [sine_wave(440.0, 0.5, 48000, 0.02)]),
        "ref_sine220_list": (48000, [np.concatenate([sine_wave(220.0, 0.25, 48000, 0.2)] * 2)]),
        "sine100_16k": (16000, [sine_wave(100.0, 1.0, 16000, 0.2)]),
        "synth16k_0": (16000, [clips16[0]]),
        "synth16k_1": (16000, [clips16[1]]),
        "synth16k_2_ragged": (16000, [clips16[2][:50001]]),
        "synth48k_0": (48000, [clips48[0]]),
        "noise": (48000, [(0.1 * rng.standard_normal(30000)).astype(np.float32)]),
        "silence": (48000, [np.zeros(5000, np.float32)]),
        "short": (48000, [sine_wave(300.0, 0.004, 48000, 0.4)]),
        "stateful_3_utts": (48000, [sine_wave(150.0, 0.3, 48000, 0.3), clips48[1][:20000],
                                    sine_wave(90.0, 0.2, 48000, 0.1)]),
    }
    return cases


def make_prosody():
    from oracle.prosody_oracle import ProsodyOracle
    arrays = {}
    for name, (sr, utts) in prosody_inputs().items():
        orc = ProsodyOracle(sample_rate=sr)
        for i, u in enumerate(utts):
            d = {}
            tags = orc.analyze_buffer(u, python_hop_loop=False, details=d)
            _, db = ProsodyOracle(sample_rate=sr).pitch_detector.run(u)  # level is state-free
            arrays[f"{name}/{i}/f0"] = d["f0"]
            arrays[f"{name}/{i}/db"] = db
            arrays[f"{name}/{i}/rms"] = np.float32(d["rms"])
            arrays[f"{name}/{i}/tags"] = np.array([tags["energy"], tags["pitch"]])
    np.savez_compressed(os.path.join(HERE, "prosody_golden.npz"), **arrays)
    print("prosody_golden.npz", len(arrays), "arrays")


def decode_setup(eot_scale):
    """Seeded weights (test dims d=128, 2 heads, 2 layers, 'confident' profile), 3 synthetic clips, oracle encoder
    states. eot_scale: <|endoftext|> embedding = eot_scale x the most frequent token's, so that hypotheses finish at
    different lengths (None: random weights, which practically never emit eot)."""
    import torch
    from janus_b200.synth import make_clips
    from janus_b200.whisper.config import mini_dims, tokenizer_spec
    from janus_b200.whisper.weights import init_random_weights
    from oracle import whisper_ref as R
    from oracle.logmel_ref import log_mel_window
    dims = mini_dims(d=128, heads=2, layers=2)
    spec = tokenizer_spec(dims)
    W = init_random_weights(dims, seed=3, profile="confident")
    clips = make_clips(3, 30.0, 16000, seed=77).numpy()
    mel = torch.stack([torch.from_numpy(log_mel_window(c, dims.n_mels)) for c in clips])
    with torch.no_grad():
        if eot_scale is not None:
            g, _ = R.greedy_decode(W, dims, spec, R.encode(W, dims, mel[:1]), 4)
            E = W["dec.tok_emb"].clone()
            E[spec.eot] = (eot_scale * E[int(g[0, 1])]).to(torch.bfloat16).float()
            W = dict(W)
            W["dec.tok_emb"] = E
        enc = R.encode(W, dims, mel)
    return dims, spec, W, enc, R


def decode_cases(n_new=16, beam=5):
    import torch
    out = {"n_new": n_new, "beam": beam, "cases": {}}
    for name, eot_scale in (("random", None), ("eot_biased", 0.9)):
        dims, spec, W, enc, R = decode_setup(eot_scale)
        with torch.no_grad():
            g_tok, g_lp, margins = R.greedy_decode(W, dims, spec, enc, n_new, return_margins=True)
            b_tok, b_lp, pools, finished = R.beam_decode(W, dims, spec, enc, n_new, beam_size=beam, return_all=True)
            stable = R.beam_stable_windows(W, dims, spec, enc, n_new, b_tok, beam_size=beam)
        case = {"greedy_tokens": g_tok.tolist(), "greedy_sum_logprob": [round(float(v), 4) for v in g_lp],
                "greedy_min_margin": [round(float(m[torch.isfinite(m)].min()), 4) for m in margins],
                "beam_tokens": b_tok.tolist(), "beam_sum_logprob": [round(float(v), 4) for v in b_lp],
                "beam_decisive": [], "finished": []}
        for b in range(enc.shape[0]):
            ranked = sorted((p[1] / max(p[2], 1) for p in pools[b]), reverse=True)
            case["beam_decisive"].append(bool(stable[b] and (len(ranked) < 2 or ranked[0] - ranked[1] > 0.05)))
            case["finished"].append([{"tokens": f[0], "score": round(float(f[1]), 4), "len": f[2]} for f in finished[b]])
        out["cases"][name] = case
    return out


def make_decode():
    out = decode_cases()
    with open(os.path.join(HERE, "decode_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("decode_golden.json", {k: len(v["beam_tokens"]) for k, v in out["cases"].items()})


if __name__ == "__main__":
    make_protocol()
    make_prosody()
    make_decode()
