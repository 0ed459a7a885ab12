"""Real-vocabulary hook: a Hugging Face `tokenizer.json` (what faster-whisper ships next to `model.bin` and loads with
`tokenizers.Tokenizer.from_file`) -> the TokenizerSpec the decoder needs + an id -> text decoder.

The reference reaches this through `WhisperModel(...).transcribe(...)` returning `segment.text`
(/root/reference/backend/services/transcriber.py:53-64). Without a vocabulary file the shared synthetic table of
`config.detokenize` is used (no Whisper vocabulary exists in this image); with one, special-token ids are looked up by
name and the non-speech suppress set is computed from the vocabulary exactly as openai-whisper /
faster-whisper's `Tokenizer.non_speech_tokens` does, instead of using the id lists recorded in config.py.
"""
from __future__ import annotations

from dataclasses import replace

from .config import TokenizerSpec, WhisperDims, tokenizer_spec

# openai-whisper tokenizer.py `non_speech_tokens`: symbols whose single-token encodings (bare or with a leading
# space) must not be sampled; the musical notes are suppressed on their first byte-level token even when they
# take several tokens.
_SYMBOLS = list("\"#()*+/:;<=>@[\\]^_`{|}~「」『』") + \
    "<< >> <<< >>> -- --- -( -[ (' (\" (( )) ((( ))) [[ ]] {{ }} ♪♪ ♪♪♪".split()
_MISC = set("♩♪♫♬♭♮♯")


class WhisperTokenizer:
    def __init__(self, tok, dims: WhisperDims | None = None) -> None:
        self.tok = tok
        self.n_vocab = tok.get_vocab_size(with_added_tokens=True)
        if dims is not None and dims.n_vocab < self.n_vocab:
            raise ValueError(f"tokenizer has {self.n_vocab} tokens, the model only {dims.n_vocab}")
        self.dims = dims

    @classmethod
    def from_file(cls, path: str, dims: WhisperDims | None = None) -> "WhisperTokenizer":
        import os
        from tokenizers import Tokenizer
        if os.path.isdir(path):
            path = os.path.join(path, "tokenizer.json")
        return cls(Tokenizer.from_file(path), dims)

    def _id(self, token: str, default: int | None = None) -> int | None:
        i = self.tok.token_to_id(token)
        return default if i is None else i

    def encode(self, text: str) -> list[int]:
        return self.tok.encode(text, add_special_tokens=False).ids

    def decode(self, ids) -> str:
        """Text tokens only (ids below <|endoftext|>), as faster-whisper's `Tokenizer.decode`."""
        eot = self._id("<|endoftext|>")
        return self.tok.decode([int(t) for t in ids if int(t) < eot], skip_special_tokens=False)

    def non_speech_tokens(self) -> tuple:
        result = set()
        for s in (" -", " '"):
            ids = self.encode(s)
            if ids:
                result.add(ids[0])
        for symbol in _SYMBOLS + sorted(_MISC):
            for ids in (self.encode(symbol), self.encode(" " + symbol)):
                if ids and (len(ids) == 1 or symbol in _MISC):
                    result.add(ids[0])
        return tuple(sorted(result))

    def spec(self) -> TokenizerSpec:
        """Special ids by NAME (so a fine-tuned vocabulary with shifted ids still works); falls back to the layout
        implied by the vocabulary size for tokens the file does not name."""
        n_vocab = self.dims.n_vocab if self.dims is not None else self.n_vocab
        base = tokenizer_spec(replace(self.dims, n_vocab=n_vocab) if self.dims is not None else
                              WhisperDims("tok", 80, 1500, 384, 6, 4, n_vocab, 448, 384, 6, 4))
        g = self._id
        ts0 = g("<|0.00|>", base.timestamp_begin)
        blank = self.encode(" ")
        return replace(
            base, eot=g("<|endoftext|>", base.eot), sot=g("<|startoftranscript|>", base.sot),
            translate=g("<|translate|>", base.translate), transcribe=g("<|transcribe|>", base.transcribe),
            sot_lm=g("<|startoflm|>", base.sot_lm), sot_prev=g("<|startofprev|>", base.sot_prev),
            no_speech=g("<|nospeech|>", g("<|nocaptions|>", base.no_speech)),
            no_timestamps=g("<|notimestamps|>", base.no_timestamps), timestamp_begin=ts0,
            lang_en=(g("<|en|>", base.lang_en) if base.lang_en is not None else None),
            blank=blank[0] if blank else base.blank, non_speech=self.non_speech_tokens(), decoder=self.decode)
