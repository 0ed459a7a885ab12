"""CPU oracle for the utterance-level loop around the Whisper model -- TEST INFRASTRUCTURE ONLY.

Restates what `WhisperModel.transcribe(audio_16k, beam_size=1, language='en')` does around encode / generate
(/root/reference/backend/services/transcriber.py:53-64 is the call site; the loop itself lives in the un-vendored
faster-whisper >= 1.0 `transcribe.py::generate_segments`, RECALLED -- PARITY UNPINNED):

  features of the WHOLE utterance once (oracle/logmel_ref.log_mel_utterance: one global max-8 clamp);
  seek = 0; while seek < content_frames (= n_frames - 1):
    window = features[:, seek : seek + 3000] zero-padded in feature space to 3000 frames; encode it
    prompt = [<|startofprev|>] + previous tokens[-(448 // 2 - 1):] (if condition_on_previous_text and there are any)
             + sot sequence
    greedy decode at temperature 0 (the temperature-fallback loop is sampling-based and stays out: the t = 0 result
    is always accepted)
    no-speech gate: skip the window (seek += segment_size, nothing emitted) iff
        no_speech_prob > no_speech_threshold (0.6) and not avg_logprob > log_prob_threshold (-1.0),
        avg_logprob = sum_logprob / (n_tokens + 1)
    segments: split at CONSECUTIVE timestamp pairs; if any exist, the window ends at the last pair (seek advances by
    that timestamp * 2 frames) unless the tokens end with a single timestamp (then seek += segment_size); with no
    consecutive pair the whole token list is one segment and seek += segment_size
    a segment whose text is empty after strip() (or whose start == end) is dropped
  text = ' '.join(segment.text.strip()).strip()        (transcriber.py:59-64)
"""
from __future__ import annotations

import torch

from oracle import whisper_ref as R
from oracle.logmel_ref import N_FRAMES, log_mel_utterance

import numpy as np


def transcribe(W, dims, spec, audio16k: np.ndarray, detok, max_new_tokens: int = 448,
               condition_on_previous_text: bool = True, no_speech_threshold: float = 0.6,
               log_prob_threshold: float = -1.0, trace: list | None = None, max_windows: int | None = None) -> str:
    """detok(list of text token ids) -> str. `trace` (optional list) receives one dict per window:
    seek, tokens, no_speech_prob, avg_logprob, skipped."""
    if len(audio16k) == 0:
        return ''
    feats = log_mel_utterance(np.asarray(audio16k, dtype=np.float32), dims.n_mels)
    content_frames = feats.shape[-1] - 1
    tb, eot = spec.timestamp_begin, spec.eot
    seek, all_tokens, texts, n_win = 0, [], [], 0
    while seek < content_frames and (max_windows is None or n_win < max_windows):
        n_win += 1
        segment_size = min(N_FRAMES, content_frames - seek)
        seg = feats[:, seek:seek + segment_size]
        seg = np.pad(seg, ((0, 0), (0, N_FRAMES - seg.shape[1])))
        prompt = list(spec.sot_sequence)
        if condition_on_previous_text and all_tokens:
            prompt = [spec.sot_prev] + all_tokens[-(spec.max_length // 2 - 1):] + prompt
        with torch.no_grad():
            enc = R.encode(W, dims, torch.from_numpy(np.ascontiguousarray(seg))[None])
            tok, sum_lp, nsp = R.greedy_decode(W, dims, spec, enc, max_new_tokens, prompt=prompt, return_no_speech=True)
        tokens = [int(t) for t in tok[0].tolist()]
        if eot in tokens:
            tokens = tokens[:tokens.index(eot)]
        tokens = tokens[:min(max_new_tokens, spec.max_length - len(prompt))]
        avg_logprob = float(sum_lp[0]) / (len(tokens) + 1)
        skip = float(nsp[0]) > no_speech_threshold and not avg_logprob > log_prob_threshold
        if trace is not None:
            trace.append({"seek": seek, "tokens": tokens, "no_speech_prob": float(nsp[0]), "avg_logprob": avg_logprob,
                          "skipped": skip, "n_prompt": len(prompt)})
        if skip:
            seek += segment_size
            continue
        is_ts = [t >= tb for t in tokens]
        single_ending = len(tokens) >= 2 and (not is_ts[-2]) and is_ts[-1]
        consecutive = [i + 1 for i in range(len(tokens) - 1) if is_ts[i] and is_ts[i + 1]]
        segs = []     # (start_pos, end_pos, tokens)
        if consecutive:
            slices = list(consecutive)
            if single_ending:
                slices.append(len(tokens))
            last = 0
            for cur in slices:
                sl = tokens[last:cur]
                segs.append((sl[0] - tb, sl[-1] - tb, sl))
                last = cur
            if single_ending:
                seek += segment_size
            else:
                seek += (tokens[last - 1] - tb) * 2       # input_stride: 2 feature frames per 0.02 s position
        else:
            stamps = [t for t in tokens if t >= tb]
            end = (stamps[-1] - tb) if (stamps and stamps[-1] != tb) else segment_size // 2
            segs.append((0, end, tokens))
            seek += segment_size
        for start, end, sl in segs:
            text = detok([t for t in sl if t < eot])
            if start == end or not text.strip():
                continue
            texts.append(text)
        all_tokens.extend(tokens)
    return ' '.join(t.strip() for t in texts).strip()
