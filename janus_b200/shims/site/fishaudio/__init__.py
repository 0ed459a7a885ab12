"""INERT `fishaudio` stub (hosted-TTS SDK of the RECEIVER side, out of scope: SURVEY.md section 8). The reference's
tests patch `FishAudio` wherever they use it; this exists only so that `from fishaudio import FishAudio` imports."""


class FishAudio:
    def __init__(self, *a, **k):
        self.tts = self

    def convert(self, *a, **k):
        raise RuntimeError("fishaudio stub: the hosted TTS is not available here")

    def stream(self, *a, **k):
        raise RuntimeError("fishaudio stub: the hosted TTS is not available here")
