"""Wire format around the packet: framing, link budget and the receiver's tag mapping,
checked against the facts the reference pins (backend/tests/test_transport_layer.py:154-335, services/engine.py)."""
import struct

import pytest

from janus_b200.protocol import JanusMode, JanusPacket
from janus_b200 import wire


def _pkts():
    return [JanusPacket("Hello world", JanusMode.SEMANTIC_VOICE, {'energy': 'Normal', 'pitch': 'High'}, None, 1234567890.0).serialize(),
            JanusPacket("x" * 300, JanusMode.TEXT_ONLY, {'energy': 'Loud', 'pitch': 'Deep'}, "Relaxed", 2.0).serialize(),
            JanusPacket("", JanusMode.MORSE_CODE, {}, None, 3.0).serialize()]


def test_tcp_frame_is_big_endian_length_prefix():
    p = _pkts()[0]
    f = wire.frame_tcp(p)
    assert f[:4] == struct.pack('>I', len(p)) and f[4:] == p          # link_simulator.py:88-92


def test_frame_many_and_stream_reassembly_with_fragmentation():
    ps = _pkts()
    stream = wire.frame_tcp_many([ps[0], None, ps[1], ps[2]])
    got, rest = wire.unframe_tcp(stream)
    assert got == ps and rest == b''
    # any split point: complete frames come out, the tail waits (recv_exact semantics, engine.py:31-52)
    for cut in (0, 1, 3, 4, 5, len(ps[0]) + 3, len(ps[0]) + 4, len(stream) - 1):
        a, tail = wire.unframe_tcp(stream[:cut])
        b, tail2 = wire.unframe_tcp(tail + stream[cut:])
        assert a + b == ps and tail2 == b''


def test_link_budget_matches_reference_constants():
    p = _pkts()[0]
    assert wire.BYTES_PER_SECOND == 37.5
    assert wire.link_delay_seconds(p) == pytest.approx(len(p) / 37.5)             # UDP
    assert wire.link_delay_seconds(p, use_tcp=True) == pytest.approx((len(p) + 4) / 37.5)


def test_receiver_tag_mapping_including_deep_quirk():
    mk = lambda e, p, o=None: JanusPacket("t", JanusMode.SEMANTIC_VOICE, {'energy': e, 'pitch': p}, o, 1.0)  # noqa: E731
    assert wire.emotion_tag(mk('Loud', 'High')) == 'Excited'
    assert wire.emotion_tag(mk('Normal', 'High')) == 'Joyful'
    assert wire.emotion_tag(mk('Loud', 'Low')) == 'Panicked'
    assert wire.emotion_tag(mk('Quiet', 'Low')) == 'Serious'
    # the sender's 'Deep' is not 'Low' on the receiver (engine.py:245-248 vs prosody.py:93): falls to Neutral
    assert wire.emotion_tag(mk('Loud', 'Deep')) == 'Neutral'
    assert wire.emotion_tag(mk('Quiet', 'Normal')) == 'Neutral'
    assert wire.emotion_tag(mk('Loud', 'High', 'Relaxed')) == 'Relaxed'
    # server path: EmotionOverride.AUTO has value "auto" != "Auto", so it is passed through as the tag
    assert wire.emotion_tag(mk('Loud', 'High', 'auto')) == 'auto'


def test_decode_many_skips_corrupt_payloads():
    ps = _pkts()
    out = wire.decode_many([ps[0], b'\xc1garbage', ps[1]])
    assert out[1] == (None, None)
    assert out[0][0].text == "Hello world" and out[0][1] == 'Joyful'
    assert out[2][0].override_emotion == "Relaxed" and out[2][1] == 'Relaxed'


def test_rich_prosody_fields_reach_the_receiver_as_numbers():
    """avg_pitch_hz / avg_energy ride in the prosody map and survive the round trip in the
    form the reference's receiver checks for (isinstance(..., (int, float)), engine.py:231-233)."""
    from janus_b200.protocol import pack_packets
    pros = [{'energy': 'Loud', 'pitch': 'High', 'avg_pitch_hz': 312.5, 'avg_energy': 0.21}]
    (pkt,) = pack_packets(["hi"], pros, JanusMode.SEMANTIC_VOICE, None, [1.0])
    back = JanusPacket.deserialize(pkt)
    assert isinstance(back.prosody['avg_pitch_hz'], float) and back.prosody['avg_pitch_hz'] == 312.5
    assert back.prosody['avg_energy'] == 0.21 and wire.emotion_tag(back) == 'Excited'
