"""`aubio` stand-in (only what /root/reference/backend/services/prosody.py:7,32-34,84 uses): `aubio.pitch('yin', buf_size,
hop_size, samplerate)` with `set_unit('Hz')`, `set_tolerance(t)`, and `detector(hop)[0]` -> f0 in Hz, a stateful
sliding buffer like the real object -- every call is one launch of the bit-exact sm_100a YIN kernel
(janus_b200/csrc/yin.cu) through the C ABI. Other aubio methods / pitch algorithms raise."""
import numpy as np

float_type = np.float32
__version__ = "0.4.9+janus_b200"


def fvec(length=1024):
    return np.zeros(int(length), dtype=np.float32)


class pitch:  # noqa: N801  (aubio's own name)
    def __init__(self, method="default", buf_size=1024, hop_size=512, samplerate=44100):
        if method not in ("yin",):
            raise RuntimeError(f"aubio shim: pitch method {method!r} is not on the Janus path (only 'yin' is)")
        from janus_b200.services.prosody import YinDetector
        self._det = YinDetector(int(buf_size), int(hop_size), int(samplerate))
        self.method, self.buf_size, self.hop_size, self.samplerate = method, int(buf_size), int(hop_size), int(samplerate)

    def set_unit(self, unit):
        self._det.set_unit(unit)

    def set_tolerance(self, tol):
        self._det.set_tolerance(tol)

    def get_tolerance(self):
        return self._det.get_tolerance()

    def set_silence(self, db):
        self._det.set_silence(db)

    def get_silence(self):
        return self._det.engine.silence_db

    def get_confidence(self):
        raise RuntimeError("aubio shim: get_confidence is not on the Janus path")

    def __call__(self, vec):
        return self._det(vec)
