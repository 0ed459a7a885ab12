"""The voice-verify step around `Transcriber.transcribe_file`.

/root/reference/backend/api/endpoints.py:52-63: the uploaded file is transcribed and accepted iff
`SequenceMatcher(None, transcript.lower().strip(), phrase.lower().strip()).ratio() >= SIMILARITY_THRESHOLD` (0.8).
Host-only string work; the FastAPI route, the upload handling and the reference-audio file stay in the reference.
"""
from __future__ import annotations

from difflib import SequenceMatcher

SIMILARITY_THRESHOLD = 0.8   # endpoints.py: SIMILARITY_THRESHOLD


def verify_transcript(transcript: str, expected_phrase: str, threshold: float = SIMILARITY_THRESHOLD):
    """-> (verified, similarity). Same normalisation and matcher as endpoints.py:58-63."""
    a = transcript.lower().strip()
    b = expected_phrase.lower().strip()
    ratio = SequenceMatcher(None, a, b).ratio()
    return ratio >= threshold, ratio


def verify_file(transcriber, file_path: str, expected_phrase: str, threshold: float = SIMILARITY_THRESHOLD) -> dict:
    """The route's result dict: {'status': 'verified'} or {'status': 'failed', 'transcript': ...}; an exception inside
    transcription becomes {'status': 'failed', 'transcript': str(e)} exactly as at endpoints.py:74-78."""
    try:
        transcript = transcriber.transcribe_file(file_path)
    except Exception as e:  # noqa: BLE001
        return {"status": "failed", "transcript": str(e)}
    ok, _ = verify_transcript(transcript, expected_phrase, threshold)
    return {"status": "verified"} if ok else {"status": "failed", "transcript": transcript}
