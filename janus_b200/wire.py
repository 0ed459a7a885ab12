"""Wire format either side of the packet: what happens to the MessagePack bytes
after `JanusPacket.serialize()` on the sender and before `JanusPacket.deserialize()` on the receiver,
for whole batches of packets at once.

Follows the reference byte for byte:
  * TCP framing: 4-byte big-endian length prefix + payload
    (/root/reference/backend/services/link_simulator.py:87-92; read back at services/engine.py:201-216)
  * UDP: the payload is the datagram, no framing (link_simulator.py:93-96)
  * the 300 bps link budget: delay = total_bytes / 37.5 s (link_simulator.py:19-20,98-99)
  * receiver-side interpretation of the prosody tags (services/engine.py:231-251), including the reference's
    quirk that the sender emits pitch 'Deep' while the receiver tests for 'Low' -- so 'Deep' maps to 'Neutral'.
Pure host code: packets are 55-500 bytes, there is nothing here for a GPU to do.
"""
from __future__ import annotations

import struct
from typing import Iterable

from .protocol import JanusPacket

BAUD_RATE = 300                      # link_simulator.py:19
BYTES_PER_SECOND = BAUD_RATE / 8     # 37.5, link_simulator.py:20
_LEN = struct.Struct('>I')


def frame_tcp(payload: bytes) -> bytes:
    """One TCP frame: big-endian uint32 length + payload (link_simulator.py:88-92)."""
    return _LEN.pack(len(payload)) + payload


def frame_tcp_many(payloads: Iterable[bytes | None]) -> bytes:
    """Concatenated frames for a batch of packets; `None` entries (no text -> nothing to send,
    engine.py:536) are skipped."""
    out = bytearray()
    for p in payloads:
        if p is None:
            continue
        out += _LEN.pack(len(p))
        out += p
    return bytes(out)


def unframe_tcp(buffer: bytes) -> tuple[list[bytes], bytes]:
    """Splits a received byte stream into complete payloads; returns (payloads, unconsumed tail).
    The stream form of `recv_exact(4)` + `recv_exact(n)` in the reference's receiver loop (engine.py:201-216)."""
    out, o, n = [], 0, len(buffer)
    while n - o >= 4:
        (ln,) = _LEN.unpack_from(buffer, o)
        if n - o - 4 < ln:
            break
        out.append(bytes(buffer[o + 4:o + 4 + ln]))
        o += 4 + ln
    return out, bytes(buffer[o:])


def link_delay_seconds(payload: bytes, use_tcp: bool = False) -> float:
    """Time the reference's 300 bps link shim spends on this packet (link_simulator.py:87-99)."""
    total = len(payload) + (4 if use_tcp else 0)
    return total / BYTES_PER_SECOND


def emotion_tag(packet: JanusPacket) -> str:
    """The receiver's prosody -> emotion mapping (engine.py:235-251), verbatim in behaviour."""
    if packet.override_emotion != "Auto":
        return packet.override_emotion
    prosody = packet.prosody
    pitch = prosody.get('pitch', 'Normal')
    energy = prosody.get('energy', 'Normal')
    if pitch == 'High' and energy == 'Loud':
        return 'Excited'
    if pitch == 'High' and energy == 'Normal':
        return 'Joyful'
    if pitch == 'Low' and energy == 'Loud':
        return 'Panicked'
    if pitch == 'Low' and energy in ('Quiet', 'Low'):
        return 'Serious'
    return 'Neutral'


def decode_many(payloads: Iterable[bytes]) -> list[tuple[JanusPacket | None, str | None]]:
    """Receiver side for a batch: (packet, emotion tag) per payload; a corrupt payload yields (None, None)
    and is skipped by the caller, like the reference's `continue` at engine.py:219-223."""
    out = []
    for data in payloads:
        try:
            pkt = JanusPacket.deserialize(data)
        except Exception:  # noqa: BLE001 -- the reference catches everything here
            out.append((None, None))
            continue
        out.append((pkt, emotion_tag(pkt)))
    return out
