"""Pins the CPU oracle (oracle/yin_aubio.c + oracle/prosody_oracle.py) against everything the
reference's own tests hold for the prosody path (backend/tests/test_input_processing.py:436-505)
and against the committed golden vectors. No GPU needed."""
import os

import numpy as np
import pytest

from janus_b200.synth import sine_wave
from oracle.prosody_oracle import AubioYinOracle, ProsodyOracle, np_mean_f32

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "prosody_golden.npz")


def test_reference_init_contract():
    e = ProsodyOracle(sample_rate=48000, hop_size=512)
    assert e.sample_rate == 48000 and e.hop_size == 512 and e.pitch_detector is not None


def test_reference_tag_facts():
    # test_input_processing.py:461-490
    assert ProsodyOracle().analyze_buffer(sine_wave(amplitude=0.02, duration=0.5))['energy'] == 'Quiet'
    assert ProsodyOracle().analyze_buffer(sine_wave(amplitude=0.1, duration=0.5))['energy'] in ('Quiet', 'Normal')
    assert ProsodyOracle().analyze_buffer(sine_wave(amplitude=0.5, duration=0.5))['energy'] in ('Normal', 'Loud')
    r = ProsodyOracle().analyze_buffer(sine_wave(frequency=440.0, duration=0.5, amplitude=0.3))
    assert r['pitch'] == 'High' and r['energy'] in ('Quiet', 'Normal', 'Loud')
    lst = [sine_wave(220.0, 0.25, amplitude=0.2)] * 2
    r = ProsodyOracle().analyze_buffer(lst)
    assert set(r) == {'energy', 'pitch'}


def test_steady_state_f0_accuracy():
    # once the 4096 buffer is full of a pure tone YIN must land within a few cents
    for sr, f in ((48000, 440.0), (16000, 100.0), (48000, 150.0)):
        d = {}
        ProsodyOracle(sample_rate=sr).analyze_buffer(sine_wave(f, 1.0, sr, 0.3), details=d)
        tail = d["f0"][12:-1]
        n = int(sr * 1.0)
        f_true = f * n / (n - 1) if False else f * (n - 1) / (n - 1)  # linspace is endpoint-inclusive
        cents = 1200 * np.abs(np.log2(tail / (f * n / (n - 1))))
        assert cents.max() < 5.0, (sr, f, cents.max())


def test_python_loop_equals_c_loop():
    x = sine_wave(180.0, 0.4, 48000, 0.2)
    d1, d2 = {}, {}
    t1 = ProsodyOracle().analyze_buffer(x, python_hop_loop=True, details=d1)
    t2 = ProsodyOracle().analyze_buffer(x, python_hop_loop=False, details=d2)
    assert t1 == t2 and np.array_equal(d1["f0"], d2["f0"])


def test_stateful_sliding_buffer():
    # the aubio object keeps its 4096 buffer across analyze_buffer calls
    a, b = sine_wave(150.0, 0.2, 48000, 0.3), sine_wave(150.0, 0.2, 48000, 0.3)
    o = ProsodyOracle()
    d1, d2 = {}, {}
    o.analyze_buffer(a, details=d1)
    o.analyze_buffer(b, details=d2)
    fresh = {}
    ProsodyOracle().analyze_buffer(b, details=fresh)
    assert not np.array_equal(d2["f0"][:7], fresh["f0"][:7])


def test_silence_gate_and_empty():
    d = {}
    r = ProsodyOracle().analyze_buffer(np.zeros(4000, np.float32), details=d)
    assert r == {'energy': 'Quiet', 'pitch': 'Normal'} and (d["f0"] == 0).all()
    r = ProsodyOracle().analyze_buffer(np.zeros(0, np.float32))
    assert r == {'energy': 'Loud', 'pitch': 'Normal'}  # nan rms falls through to 'Loud' in the reference


def test_numpy_pairwise_restatement():
    rng = np.random.default_rng(0)
    for n in list(range(1, 300)) + [1000, 2813, 4097, 100001]:
        a = rng.uniform(50, 400, n).astype(np.float32)
        assert np.mean(list(a)) == np_mean_f32(a), n


def test_golden_vectors():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    from make_golden import prosody_inputs
    g = np.load(GOLDEN)
    for name, (sr, utts) in prosody_inputs().items():
        o = ProsodyOracle(sample_rate=sr)
        for i, u in enumerate(utts):
            d = {}
            tags = o.analyze_buffer(u, python_hop_loop=False, details=d)
            assert np.array_equal(d["f0"], g[f"{name}/{i}/f0"]), name
            assert [tags["energy"], tags["pitch"]] == list(g[f"{name}/{i}/tags"]), name


def test_detector_rejects_wrong_hop():
    with pytest.raises(ValueError):
        AubioYinOracle()(np.zeros(100, np.float32))
