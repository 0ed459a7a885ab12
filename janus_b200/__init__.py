"""janus_b200 -- B200-native (sm_100a) implementation of the sender-side semantic-encode path of
akshatvasisht/janus: Whisper STT + YIN/RMS prosody -> MessagePack JanusPacket.

Host code is Python (PyTorch tensors for device memory/streams only); the arithmetic lives in
hand-written CUDA kernels in libjanus_b200.so, reached through the C ABI in include/janus_b200.h.
There is no CPU fallback: importing the service classes without the built library, or running
them without an sm_100 device, raises.
"""
__version__ = "0.1.0"

import os as _os

# The batched engine runs one encoder stream plus up to ~16 decode-lane streams concurrently; with the
# default of 8 hardware work queues independent streams would be serialised behind each other.
# Must be set before the CUDA context is created (harmless otherwise).
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
