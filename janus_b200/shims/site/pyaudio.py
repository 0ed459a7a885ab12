"""INERT `pyaudio` stub: lets the reference's modules import on a machine without PortAudio (the reference's own tests
replace AudioService with a mock: backend/tests/conftest.py:120-154). Opening a stream raises OSError, which
backend/services/audio_io.py already treats as "no audio hardware". Not part of the hot path."""
paInt16 = 8
paFloat32 = 1
paContinue = 0


class PyAudio:
    def open(self, *a, **k):
        raise OSError("pyaudio stub: no audio hardware on this machine")

    def get_sample_size(self, fmt):
        return 2

    def get_default_input_device_info(self):
        raise OSError("pyaudio stub: no audio hardware on this machine")

    def terminate(self):
        pass
