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


def test_voice_verify_similarity_rule():
    """endpoints.py:58-63 on host strings (the GPU leg is tests/test_gpu_transcribe.py)."""
    from janus_b200.services.verify import SIMILARITY_THRESHOLD, verify_file, verify_transcript
    phrase = "The quick brown fox jumps over the lazy dog"
    assert SIMILARITY_THRESHOLD == 0.8
    assert verify_transcript("  the quick brown fox jumps over the lazy dog ", phrase) == (True, 1.0)
    ok, r = verify_transcript("the quick brown fox jumped over a lazy dog", phrase)
    assert ok and 0.8 <= r < 1.0
    ok, r = verify_transcript("hello there", phrase)
    assert not ok and r < 0.8

    class T:
        def __init__(self, out): self.out = out
        def transcribe_file(self, p):
            if isinstance(self.out, Exception):
                raise self.out
            return self.out
    assert verify_file(T(phrase.upper()), "x.wav", phrase) == {"status": "verified"}
    assert verify_file(T("nope"), "x.wav", phrase) == {"status": "failed", "transcript": "nope"}
    assert verify_file(T(ValueError("bad wav")), "x.wav", phrase) == {"status": "failed", "transcript": "bad wav"}


def test_split_window_segments_rules():
    """The segment / seek rule of the utterance loop (services/transcriber.py; faster-whisper generate_segments)."""
    from janus_b200.services.transcriber import split_window_segments
    tb = 1000
    # <0.00> a b <1.00><1.00> c <2.00>  -> two segments, ends in a single timestamp: whole window consumed
    segs, adv = split_window_segments([tb, 1, 2, tb + 50, tb + 50, 3, tb + 100], tb, 3000)
    assert [s[:2] for s in segs] == [(0, 50), (50, 100)] and adv == 3000
    assert segs[0][2] == [tb, 1, 2, tb + 50] and segs[1][2] == [tb + 50, 3, tb + 100]
    # ends with a consecutive pair -> seek advances to that timestamp (2 frames per 0.02 s position)
    segs, adv = split_window_segments([tb, 1, tb + 40, tb + 40, 2, tb + 90, tb + 90], tb, 3000)
    assert [s[:2] for s in segs] == [(0, 40), (40, 90)] and adv == 180
    # tokens after the last pair are dropped from this window and re-decoded from the new seek
    segs, adv = split_window_segments([tb, 1, tb + 40, tb + 40, 2, 3], tb, 3000)
    assert [s[:2] for s in segs] == [(0, 40)] and adv == 80
    # no consecutive pair: one segment, duration from the last timestamp, whole window consumed
    segs, adv = split_window_segments([tb + 5, 1, 2, 3], tb, 1200)
    assert segs == [(0, 5, [tb + 5, 1, 2, 3])] and adv == 1200
    segs, adv = split_window_segments([1, 2], tb, 1200)
    assert segs == [(0, 600, [1, 2])] and adv == 1200
    assert split_window_segments([], tb, 700) == ([(0, 350, [])], 700)
