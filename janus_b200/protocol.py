"""JanusPacket wire contract (sender side), byte-identical to the reference.

Mirrors /root/reference/backend/common/protocol.py: `JanusMode` (:15-21), `JanusPacket`
constructor (:30-55), `to_dict` (:57-76), `from_dict` (:78-95), `serialize` (:97-107),
`deserialize` (:109-121). The wire form is a MessagePack map with insertion-ordered compact keys
't' (text), 'm' (mode int), 'p' (prosody map), 'ts' (float64 timestamp) and -- only when the
override differs from the literal "Auto" -- 'o'.

`pack_packets` is the batched builder used after a multi-stream GPU step: one packer object,
no per-packet dict churn, same bytes.
"""
from __future__ import annotations

import enum
import time
from typing import Iterable, Optional, Sequence

import msgpack

_AUTO = "Auto"


class JanusMode(enum.IntEnum):
    SEMANTIC_VOICE = 0  # text + prosody
    TEXT_ONLY = 1       # text, receiver uses its default voice
    MORSE_CODE = 2      # receiver synthesises morse locally


class JanusPacket:
    __slots__ = ("text", "mode", "prosody", "override_emotion", "timestamp")

    def __init__(self, text: str, mode: JanusMode, prosody: dict[str, str],
                 override_emotion: Optional[str] = None, timestamp: Optional[float] = None) -> None:
        self.text = text
        self.mode = mode
        self.prosody = prosody
        self.override_emotion = _AUTO if override_emotion is None else override_emotion
        self.timestamp = time.time() if timestamp is None else timestamp

    # -- wire form ---------------------------------------------------------------------------
    def to_dict(self) -> dict:
        wire = {'t': self.text, 'm': int(self.mode), 'p': self.prosody, 'ts': self.timestamp}
        if self.override_emotion != _AUTO:
            wire['o'] = self.override_emotion
        return wire

    def serialize(self) -> bytes:
        return msgpack.packb(self.to_dict(), use_bin_type=True)

    @classmethod
    def from_dict(cls, data: dict) -> "JanusPacket":
        return cls(
            data.get('t', ''),
            JanusMode(data.get('m', 0)),
            data.get('p', {}),
            data.get('o', _AUTO),
            data.get('ts', time.time()),
        )

    @classmethod
    def deserialize(cls, payload_bytes: bytes) -> "JanusPacket":
        return cls.from_dict(msgpack.unpackb(payload_bytes, raw=False))

    def __repr__(self) -> str:  # debugging aid only
        return (f"JanusPacket(text={self.text!r}, mode={self.mode!r}, prosody={self.prosody!r}, "
                f"override_emotion={self.override_emotion!r}, timestamp={self.timestamp!r})")


def pack_packets(texts: Sequence[str], prosodies: Iterable[dict[str, str]],
                 mode: JanusMode = JanusMode.SEMANTIC_VOICE, override_emotion: Optional[str] = None,
                 timestamps: Optional[Sequence[float]] = None,
                 skip_empty_text: bool = True) -> list[Optional[bytes]]:
    """Builds one packet per stream after a batched step. Streams whose text is empty after
    strip() produce None when `skip_empty_text` (the reference does not transmit them:
    /root/reference/backend/services/engine.py:536)."""
    packer = msgpack.Packer(use_bin_type=True)
    override = _AUTO if override_emotion is None else override_emotion
    m = int(mode)
    now = time.time()
    out: list[Optional[bytes]] = []
    for i, (text, pros) in enumerate(zip(texts, prosodies)):
        if skip_empty_text and not text.strip():
            out.append(None)
            continue
        wire = {'t': text, 'm': m, 'p': pros, 'ts': now if timestamps is None else timestamps[i]}
        if override != _AUTO:
            wire['o'] = override
        out.append(packer.pack(wire))
    return out
