"""INERT stub of fishaudio.types (see fishaudio/__init__.py)."""


class ReferenceAudio:
    def __init__(self, audio=b"", text=""):
        self.audio, self.text = audio, text
