"""Utterance segmentation against a literal restatement of the reference loop."""
from collections import deque

import numpy as np

from janus_b200.segmenter import (CHUNK_SIZE, MIN_UTTERANCE_SAMPLES, MultiSessionSegmenter, UtteranceSegmenter,
                                  energy_gate)


def _reference_loop(chunks, speech_flags):
    """engine.py:438-506, streaming mode, VAD decisions given; returns the list of emitted buffers."""
    audio_buffer, pre_roll, silence_counter, out = [], deque(maxlen=10), 0, []
    for chunk, is_speech in zip(chunks, speech_flags):
        trigger = False
        if is_speech:
            if len(audio_buffer) == 0:
                audio_buffer.extend(list(pre_roll))
            audio_buffer.append(chunk)
            silence_counter = 0
        else:
            silence_counter += 1
            if len(audio_buffer) > 0:
                audio_buffer.append(chunk)
            else:
                pre_roll.append(chunk)
            if silence_counter > 15:
                trigger = True
        if trigger and len(audio_buffer) > 0:
            combined = np.concatenate(audio_buffer)
            audio_buffer, silence_counter = [], 0
            if len(combined) < 1536 * 6:
                continue
            out.append(combined)
    return out


def _pattern(seed, n=400):
    rng = np.random.default_rng(seed)
    flags, chunks = [], []
    state = False
    while len(flags) < n:
        run = int(rng.integers(1, 40)) if not state else int(rng.integers(1, 25))
        flags += [state] * run
        state = not state
    flags = flags[:n]
    for i, f in enumerate(flags):
        amp = 0.3 if f else 0.0005
        chunks.append((amp * rng.standard_normal(CHUNK_SIZE)).astype(np.float32) + np.float32(i) * 0)
    return chunks, flags


def test_matches_reference_state_machine():
    for seed in range(5):
        chunks, flags = _pattern(seed)
        seg = UtteranceSegmenter()
        got = [u for u in (seg.push(c, f) for c, f in zip(chunks, flags)) if u is not None]
        ref = _reference_loop(chunks, flags)
        assert len(got) == len(ref) and len(ref) > 0
        for a, b in zip(got, ref):
            assert np.array_equal(a, b)


def test_pre_roll_trigger_point_and_short_skip():
    z = np.zeros(CHUNK_SIZE, np.float32)
    s = np.ones(CHUNK_SIZE, np.float32)
    seg = UtteranceSegmenter()
    for _ in range(12):
        assert seg.push(z, False) is None          # only the last 10 are kept as pre-roll
    for _ in range(3):
        assert seg.push(s, True) is None
    outs = [seg.push(z, False) for _ in range(16)]
    assert all(o is None for o in outs[:15]) and outs[15] is not None   # ends on the 16th silent chunk (> 15)
    assert len(outs[15]) == (10 + 3 + 16) * CHUNK_SIZE
    # a buffer shorter than 1536*6 samples cannot arise through the gate (16 trailing chunks), only via flush
    seg2 = UtteranceSegmenter()
    seg2.push(s, True)
    assert seg2.flush() is None and seg2.skipped_short == 1
    for _ in range(6):
        seg2.push(s, True)
    assert len(seg2.flush()) == 6 * CHUNK_SIZE == MIN_UTTERANCE_SAMPLES


def test_multi_session_interleaved_equals_per_session():
    streams = {sid: _pattern(10 + sid, n=200) for sid in range(4)}
    ms = MultiSessionSegmenter(is_speech=energy_gate(-40.0))
    for i in range(200):
        for sid, (chunks, flags) in streams.items():
            ms.push(sid, chunks[i])                 # decisions from the energy gate: 0.3 rms vs 0.0005 rms
    got = ms.take()
    assert ms.take() == []
    for sid, (chunks, flags) in streams.items():
        ref = _reference_loop(chunks, flags)
        mine = [u for s, u in got if s == sid]
        assert len(mine) == len(ref)
        for a, b in zip(mine, ref):
            assert np.array_equal(a, b)


def test_push_tick_equals_per_session_push():
    from janus_b200.segmenter import MultiSessionSegmenter
    rng = np.random.default_rng(5)
    a, b = MultiSessionSegmenter(), MultiSessionSegmenter()
    for t in range(80):
        chunks = rng.standard_normal((4, 1536)).astype(np.float32)
        speech = rng.random(4) < (0.7 if (t // 20) % 2 == 0 else 0.02)
        a.push_tick(range(4), chunks, speech)
        for s in range(4):
            b.push(s, chunks[s], bool(speech[s]))
    ra, rb = a.take(), b.take()
    assert len(ra) == len(rb) > 0
    for (sa, ua), (sb, ub) in zip(ra, rb):
        assert sa == sb and np.array_equal(ua, ub)
