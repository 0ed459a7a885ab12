"""GPU speech gate for many sessions (janus_b200/vad.py, vad_gate_kernel) against its CPU restatement over the
aubio-equivalent YIN oracle, chunk for chunk with per-session state, and through the reference's utterance state
machine (janus_b200/segmenter.py; /root/reference/backend/services/engine.py:471-506)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _session_audio(n_sessions, n_chunks, seed):
    """48 kHz streams made of talk spurts (harmonic tone + noise) and pauses (faint noise / digital silence)."""
    from janus_b200.synth import make_clips
    rng = np.random.default_rng(seed)
    N = n_chunks * 1536
    tone = make_clips(n_sessions, N / 48000.0, 48000, seed=seed).numpy()
    out = np.zeros((n_sessions, N), np.float32)
    for s in range(n_sessions):
        t = 0
        talking = bool(rng.integers(2))
        while t < N:
            d = int(rng.uniform(0.3, 1.2) * 48000)
            if talking:
                out[s, t:t + d] = tone[s, t:t + d]
            elif rng.random() < 0.5:
                out[s, t:t + d] = (1e-4 * rng.standard_normal(min(d, N - t))).astype(np.float32)
            t += d
            talking = not talking
    return out


def test_gate_flags_and_utterances_match_the_oracle():
    from janus_b200.segmenter import MultiSessionSegmenter
    from janus_b200.vad import GpuVadGate
    from oracle.prosody_oracle import AubioYinOracle, vad_gate_ref
    S, T = 6, 120
    audio = _session_audio(S, T, seed=17)
    gate = GpuVadGate(S, level_db=-40.0, min_voiced=1)
    seg_gpu, seg_ref = MultiSessionSegmenter(), MultiSessionSegmenter()
    dets = []
    for _ in range(S):
        d = AubioYinOracle(4096, 512, 48000)
        d.set_unit('Hz'); d.set_tolerance(0.8)
        dets.append(d)
    n_speech = n_total = 0
    for t in range(T):
        chunks = audio[:, t * 1536:(t + 1) * 1536]
        flags = gate.step(chunks)
        ref = np.zeros(S, bool)
        for s in range(S):
            f0, db = dets[s].run(chunks[s])            # stateful: the detector keeps its sliding buffer
            ref[s] = vad_gate_ref(f0, db, 3, -40.0, 1)[0]
        assert np.array_equal(flags, ref), (t, flags, ref)
        n_speech += int(flags.sum()); n_total += S
        seg_gpu.push_tick(range(S), chunks, flags)
        seg_ref.push_tick(range(S), chunks, ref)
    assert 0.2 < n_speech / n_total < 0.9          # both outcomes occur
    a, b = seg_gpu.take(), seg_ref.take()
    assert len(a) == len(b) >= S // 2
    for (sa, ua), (sb, ub) in zip(a, b):
        assert sa == sb and np.array_equal(ua, ub)


def test_gate_min_voiced_and_level_are_inputs():
    """level_db and min_voiced are inputs. `voiced` means f0 > 0, i.e. the hop passed aubio's -50 dB silence gate and
    YIN returned a period -- with aubio's fallback to the global minimum that is true for any non-silent hop, noise
    included, so min_voiced adds the -50 dB floor, not a periodicity test."""
    from janus_b200.synth import sine_wave
    from janus_b200.vad import GpuVadGate
    rng = np.random.default_rng(3)
    tone = sine_wave(200.0, 0.5, 48000, 0.3)[:1536 * 8].reshape(8, 1536)
    faint = (1e-3 * rng.standard_normal((8, 1536))).astype(np.float32)       # about -60 dB: below aubio's gate
    g1 = GpuVadGate(2, level_db=-100.0, min_voiced=1)
    g0 = GpuVadGate(2, level_db=-100.0, min_voiced=0)
    quiet = GpuVadGate(2, level_db=-5.0, min_voiced=0)
    f1 = np.array([g1.step(np.stack([tone[i], faint[i]])) for i in range(8)])
    f0 = np.array([g0.step(np.stack([tone[i], faint[i]])) for i in range(8)])
    fq = np.array([quiet.step(np.stack([tone[i], faint[i]])) for i in range(8)])
    assert f1[:, 0].all() and f0[:, 0].all()   # the tone is speech under both settings
    assert f0[:, 1].all()                       # level -100 dB alone lets the faint noise through ...
    assert not f1[:, 1].any()                   # ... requiring a voiced hop applies aubio's -50 dB silence gate
    assert not fq.any()                         # nothing is louder than -5 dB
