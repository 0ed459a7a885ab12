"""Packet builder parity: janus_b200.protocol vs golden bytes produced by the reference's own
protocol.py (tests/golden/make_golden.py), plus the facts the reference's
backend/tests/test_transport_layer.py:26-147 pins (round trip, compact keys, 'o' rule, garbage)."""
import enum
import json
import os
import time

import msgpack
import pytest

from janus_b200.protocol import JanusMode, JanusPacket, pack_packets

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "protocol_golden.json")


def _cases():
    with open(GOLDEN) as f:
        return json.load(f)


class _StrEnum(str, enum.Enum):
    AUTO = "auto"


@pytest.mark.parametrize("case", _cases(), ids=lambda c: c["text"][:12] or "empty")
def test_golden_bytes(case):
    ov = case["override"]
    if case["override_is_str_enum"]:
        ov = _StrEnum(ov)
    pkt = JanusPacket(case["text"], JanusMode(case["mode"]), case["prosody"], ov, case["ts"])
    assert pkt.serialize().hex() == case["hex"]
    back = JanusPacket.deserialize(bytes.fromhex(case["hex"]))
    assert back.text == case["text"] and int(back.mode) == case["mode"]
    assert back.prosody == case["prosody"] and back.timestamp == case["ts"]


def test_known_hex_from_survey():
    b = JanusPacket("Hello world", JanusMode.SEMANTIC_VOICE, {'energy': 'Normal', 'pitch': 'High'},
                    None, 1234567890.0).serialize()
    assert b.hex() == ("84a174ab48656c6c6f20776f726c64a16d00a17082a6656e65726779a64e6f726d616c"
                       "a57069746368a448696768a27473cb41d26580b4800000")


def test_roundtrip_and_keys():
    p = JanusPacket("Test message", JanusMode.TEXT_ONLY, {'energy': 'Loud', 'pitch': 'Deep'},
                    override_emotion="Excited", timestamp=1234567890.0)
    d = p.to_dict()
    assert list(d.keys()) == ['t', 'm', 'p', 'ts', 'o'] and d['m'] == 1
    q = JanusPacket.deserialize(p.serialize())
    assert (q.text, q.mode, q.prosody, q.override_emotion, q.timestamp) == \
        ("Test message", JanusMode.TEXT_ONLY, {'energy': 'Loud', 'pitch': 'Deep'}, "Excited", 1234567890.0)


def test_override_auto_omitted():
    assert 'o' not in JanusPacket("x", JanusMode.SEMANTIC_VOICE, {}, "Auto").to_dict()
    assert 'o' not in JanusPacket("x", JanusMode.SEMANTIC_VOICE, {}).to_dict()
    assert JanusPacket("x", JanusMode.SEMANTIC_VOICE, {}, "auto").to_dict()['o'] == "auto"


def test_default_timestamp_and_from_dict_defaults():
    t0 = time.time()
    p = JanusPacket("x", JanusMode.MORSE_CODE, {})
    assert t0 <= p.timestamp <= time.time()
    q = JanusPacket.from_dict({})
    assert q.text == '' and q.mode == JanusMode.SEMANTIC_VOICE and q.prosody == {} and q.override_emotion == "Auto"


def test_garbage_raises():
    with pytest.raises(Exception):
        JanusPacket.deserialize(b"\xc1\xff\x00garbage")


def test_pack_packets_matches_serialize():
    texts = ["hello", "", "  ", "world " * 20]
    pros = [{'energy': 'Quiet', 'pitch': 'Deep'}] * 4
    ts = [1.0, 2.0, 3.0, 4.0]
    out = pack_packets(texts, pros, JanusMode.SEMANTIC_VOICE, "auto", ts)
    assert out[1] is None and out[2] is None
    for i in (0, 3):
        assert out[i] == JanusPacket(texts[i], JanusMode.SEMANTIC_VOICE, pros[i], "auto", ts[i]).serialize()
        assert msgpack.unpackb(out[i])['t'] == texts[i]


def _reference_protocol():
    """The reference's own protocol.py (mounted at /root/reference or staged, unmodified, under baseline/_ref)."""
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for base in ("/root/reference", os.path.join(root, "baseline", "_ref")):
        path = os.path.join(base, "backend", "common", "protocol.py")
        if os.path.exists(path):
            spec = importlib.util.spec_from_file_location("ref_protocol_fuzz", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    pytest.skip("reference protocol.py not available on this machine")


def test_fuzz_against_the_reference_module():
    """Property test (hypothesis): for arbitrary unicode text, every mode, arbitrary prosody maps (incl. the numeric
    avg_pitch_hz / avg_energy extension), overrides ('Auto' / None / any string / a str-Enum as the server passes) and
    float timestamps, `JanusPacket.serialize()` and `pack_packets` emit exactly the reference's bytes, and the
    reference's deserialize reads them back."""
    from hypothesis import given, settings, strategies as st
    ref = _reference_protocol()
    text = st.text(max_size=400)
    tags = st.dictionaries(st.sampled_from(["energy", "pitch", "avg_pitch_hz", "avg_energy", "x"]),
                           st.one_of(st.sampled_from(["Quiet", "Normal", "Loud", "Deep", "High"]),
                                     st.floats(allow_nan=False, allow_infinity=False, width=32)), max_size=4)
    override = st.one_of(st.none(), st.just("Auto"), st.just("auto"), st.text(max_size=12), st.just(_StrEnum.AUTO))
    ts = st.floats(min_value=0.0, max_value=4e9, allow_nan=False)

    @settings(max_examples=300, deadline=None)
    @given(text, st.sampled_from([0, 1, 2]), tags, override, ts)
    def check(t, m, p, o, stamp):
        want = ref.JanusPacket(t, ref.JanusMode(m), p, o, stamp).serialize()
        assert JanusPacket(t, JanusMode(m), p, o, stamp).serialize() == want
        got = pack_packets([t], [p], JanusMode(m), o, [stamp], skip_empty_text=False)[0]
        assert got == want
        back = ref.JanusPacket.deserialize(got)
        assert back.text == t and int(back.mode) == m and back.prosody == p and back.timestamp == stamp

    check()
