"""Module-level stand-ins for the reference's un-vendored third-party packages, backed by libjanus_b200.so.

`janus_b200/shims/site/` is a directory meant to be PREPENDED to PYTHONPATH of an unmodified Janus checkout:

    PYTHONPATH=/path/to/janus_b200/shims/site:/path/to/repo  python -m backend.server

`import aubio` / `from faster_whisper import WhisperModel` inside the reference's own
backend/services/prosody.py and backend/services/transcriber.py then resolve to the B200 kernels -- the second
integration route of INTEGRATION.md (the first replaces the two service modules). `pyaudio` and `fishaudio` there are
INERT stubs for machines without audio hardware / the hosted TTS SDK (the reference's own tests mock both); they exist
so that the reference's test-suite imports, and are not part of the path.
"""
import os

SITE_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "site")
