"""Single-utterance latency through the actual drop-in calls -- the only way the reference ever uses the path
(/root/reference/backend/services/engine.py:510-531: one 1-10 s utterance at a time, `Transcriber()` = base.en,
`ProsodyExtractor()` at 48 kHz). p50 / p90 over N calls of `transcribe_buffer` and `analyze_buffer` on 2 s and 10 s
buffers, next to the CPU oracle port on the same buffers.

Random-init weights never emit <|endoftext|>, so the token budget is capped at 32 new tokens per window on both sides
(a 2-10 s utterance is ~10-30 tokens); one window per call.

    python scripts/latency_transcriber.py > profiles/r02_latency_single_utterance.json
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N_CALLS = int(os.environ.get("CALLS", "50"))
N_NEW = 32


def pct(v, q):
    return float(np.percentile(np.asarray(v), q))


def main():
    import torch
    import janus_b200.services.transcriber as T
    from janus_b200.services.prosody import ProsodyExtractor
    from janus_b200.synth import make_clips
    from janus_b200.whisper.config import detokenize
    from janus_b200.whisper.weights import init_random_weights
    from oracle.prosody_oracle import ProsodyOracle
    from oracle.transcribe_ref import transcribe
    T.MAX_NEW_TOKENS = N_NEW
    tr = T.Transcriber()                 # base.en, as engine.py:412
    tr.max_windows = 1
    pe = ProsodyExtractor()
    W = init_random_weights(tr.dims, seed=0, profile="confident")
    torch.set_num_threads(os.cpu_count() or 1)
    out = {"model": tr.dims.name, "max_new_tokens": N_NEW, "calls": N_CALLS, "cores": os.cpu_count(), "cases": {}}
    for seconds in (2.0, 10.0):
        x48 = make_clips(1, seconds, 48000, seed=71).numpy()[0]
        for _ in range(3):
            tr.transcribe_buffer(x48); pe.analyze_buffer(x48)
        stt, pro = [], []
        for _ in range(N_CALLS):
            t0 = time.perf_counter(); text = tr.transcribe_buffer(x48); stt.append(time.perf_counter() - t0)
            t0 = time.perf_counter(); tags = pe.analyze_buffer(x48); pro.append(time.perf_counter() - t0)
        n_tok = len(tr.last_trace[0]["tokens"])
        # CPU port, same buffer (3 calls: it is slow)
        cpu_stt, cpu_pro = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            ref = transcribe(W, tr.dims, tr.spec, x48[::3], lambda ids: detokenize(ids, tr.spec), max_new_tokens=N_NEW,
                             max_windows=1)
            cpu_stt.append(time.perf_counter() - t0)
            t0 = time.perf_counter(); otags = ProsodyOracle().analyze_buffer(x48, python_hop_loop=True)
            cpu_pro.append(time.perf_counter() - t0)
        out["cases"][f"{seconds:g}s"] = {
            "tokens_decoded": n_tok, "text_equals_cpu_port": text == ref,
            "stt_ms": {"p50": 1e3 * pct(stt, 50), "p90": 1e3 * pct(stt, 90), "min": 1e3 * min(stt)},
            "prosody_ms": {"p50": 1e3 * pct(pro, 50), "p90": 1e3 * pct(pro, 90), "min": 1e3 * min(pro)},
            "cpu_port_stt_ms": {"p50": 1e3 * pct(cpu_stt, 50)}, "cpu_port_prosody_ms": {"p50": 1e3 * pct(cpu_pro, 50)}}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
