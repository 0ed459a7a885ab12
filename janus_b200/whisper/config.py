"""Whisper architecture descriptions (dimensions + special-token layout).

These are the public OpenAI Whisper model shapes that faster-whisper loads by name
(/root/reference/backend/services/transcriber.py:23-27 passes `model_size`, default 'base.en').
No tokenizer vocabulary file exists offline, so everything token-dependent (special ids, the
non-speech suppress list, the blank token) is explicit data here and is an INPUT to both the CPU
oracle and the CUDA decoder.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace


@dataclass(frozen=True)
class WhisperDims:
    name: str
    n_mels: int
    n_audio_ctx: int
    n_audio_state: int
    n_audio_head: int
    n_audio_layer: int
    n_vocab: int
    n_text_ctx: int
    n_text_state: int
    n_text_head: int
    n_text_layer: int

    @property
    def d(self) -> int:
        return self.n_audio_state

    @property
    def head_dim(self) -> int:
        return self.n_audio_state // self.n_audio_head


def _dims(name, n_mels, d, heads, layers, vocab) -> WhisperDims:
    return WhisperDims(name, n_mels, 1500, d, heads, layers, vocab, 448, d, heads, layers)


PRESETS = {
    "tiny": _dims("tiny", 80, 384, 6, 4, 51865),
    "tiny.en": _dims("tiny.en", 80, 384, 6, 4, 51864),
    "base": _dims("base", 80, 512, 8, 6, 51865),
    "base.en": _dims("base.en", 80, 512, 8, 6, 51864),
    "small": _dims("small", 80, 768, 12, 12, 51865),
    "small.en": _dims("small.en", 80, 768, 12, 12, 51864),
    "medium": _dims("medium", 80, 1024, 16, 24, 51865),
    "large-v3": _dims("large-v3", 128, 1280, 20, 32, 51866),
}


def mini_dims(d: int = 128, heads: int = 2, layers: int = 2, n_mels: int = 80, vocab: int = 51865,
              n_audio_ctx: int = 1500) -> WhisperDims:
    """A shrunken architecture for fast parity tests (same code paths, same vocab layout)."""
    return WhisperDims(f"test-d{d}-l{layers}", n_mels, n_audio_ctx, d, heads, layers, vocab, 448, d, heads, layers)


# The non-speech suppress list shipped in the public multilingual Whisper generation configs
# (what faster-whisper's `suppress_tokens=[-1]` expands to via tokenizer.non_speech_tokens).
_NON_SPEECH_MULTI = (
    1, 2, 7, 8, 9, 10, 14, 25, 26, 27, 28, 29, 31, 58, 59, 60, 61, 62, 63, 90, 91, 92, 93, 359, 503, 522,
    542, 873, 893, 902, 918, 922, 931, 1350, 1853, 1982, 2460, 2627, 3246, 3253, 3268, 3536, 3846, 3961,
    4183, 4667, 6585, 6647, 7273, 9061, 9383, 10428, 10929, 11938, 12033, 12331, 12562, 13793, 14157,
    14635, 15265, 15618, 16553, 16604, 18362, 18956, 20075, 21675, 22520, 26130, 26161, 26435, 28279,
    29464, 31650, 32302, 32470, 36865, 42863, 47425, 49870, 50254)
_NON_SPEECH_EN = (
    1, 2, 7, 8, 9, 10, 14, 25, 26, 27, 28, 29, 31, 58, 59, 60, 61, 62, 63, 90, 91, 92, 93, 357, 366, 438,
    532, 685, 705, 796, 930, 1058, 1220, 1267, 1279, 1303, 1343, 1377, 1391, 1635, 1782, 1875, 2162,
    2361, 2488, 3467, 4008, 4211, 4600, 4808, 5299, 5855, 6329, 7203, 9609, 9959, 10563, 10786, 11420,
    11709, 11907, 13163, 13697, 13700, 14808, 15306, 16410, 16791, 17992, 19203, 19510, 20724, 22305,
    22935, 27007, 30109, 30420, 33409, 34949, 40283, 40493, 40549, 47282, 49146)


@dataclass(frozen=True)
class TokenizerSpec:
    """Special-token layout + decode-time filters (faster-whisper transcribe() defaults)."""
    n_vocab: int
    eot: int
    sot: int
    translate: int
    transcribe: int
    sot_lm: int
    sot_prev: int
    no_speech: int
    no_timestamps: int
    timestamp_begin: int
    lang_en: int | None           # None for English-only models
    blank: int = 220              # the " " token (suppress_blank=True)
    non_speech: tuple = ()
    max_initial_timestamp_index: int = 50
    max_length: int = 448
    # id -> text for the text tokens of one segment (a real vocabulary: whisper/tokenizer.py); None = the shared
    # synthetic table below (no Whisper vocabulary exists offline)
    decoder: object = field(default=None, compare=False, repr=False)

    @property
    def sot_sequence(self) -> tuple:
        # language='en', task transcribe, timestamps on (without_timestamps=False)
        if self.lang_en is None:
            return (self.sot,)
        return (self.sot, self.lang_en, self.transcribe)

    @property
    def suppress_ids(self) -> tuple:
        # get_suppressed_tokens(): non-speech set + [transcribe, translate, sot, sot_prev, sot_lm];
        # the timestamp rules additionally suppress <|notimestamps|>; no_speech is never emitted.
        ids = set(self.non_speech) | {self.transcribe, self.translate, self.sot, self.sot_prev,
                                      self.sot_lm, self.no_speech, self.no_timestamps}
        return tuple(sorted(i for i in ids if 0 <= i < self.n_vocab))


def tokenizer_spec(dims: WhisperDims) -> TokenizerSpec:
    v = dims.n_vocab
    if v == 51864:  # English-only (gpt2 vocabulary)
        return TokenizerSpec(v, eot=50256, sot=50257, translate=50357, transcribe=50358, sot_lm=50359,
                             sot_prev=50360, no_speech=50361, no_timestamps=50362, timestamp_begin=50363,
                             lang_en=None, non_speech=_NON_SPEECH_EN)
    n_lang = v - 51865 + 99  # 99 languages (51865) or 100 (large-v3, 51866)
    base = 50259 + n_lang
    return TokenizerSpec(v, eot=50257, sot=50258, translate=base, transcribe=base + 1, sot_lm=base + 2,
                         sot_prev=base + 3, no_speech=base + 4, no_timestamps=base + 5,
                         timestamp_begin=base + 6, lang_en=50259, non_speech=_NON_SPEECH_MULTI)


def detokenize(ids, spec: TokenizerSpec) -> str:
    """Shared deterministic id -> string table. No Whisper vocabulary exists offline (SURVEY.md
    section 3.5), so text tokens map to a fixed synthetic word list; special/timestamp tokens map
    to nothing, exactly as a real tokenizer's decode() drops them. With a real vocabulary loaded
    (`spec.decoder`, whisper/tokenizer.py) that tokenizer's own decode is used instead."""
    if spec.decoder is not None:
        return spec.decoder(ids)
    words = []
    for t in ids:
        t = int(t)
        if t >= spec.eot:
            continue
        words.append(f" w{t}")
    return "".join(words)
