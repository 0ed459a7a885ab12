"""The reference's OWN tests and service classes, unmodified, over janus_b200/shims/site on a real B200 (SURVEY.md
section 4 "useful trick", row A10): `import aubio` / `from faster_whisper import WhisperModel` inside the reference's
backend/services/prosody.py and transcriber.py resolve to the CUDA kernels behind the C ABI.

The reference tree is read from /root/reference when it is mounted, else from baseline/_ref (an unmodified copy staged
by scripts/stage_reference_tests.py; git-ignored, travels with the gpurun snapshot). Neither present -> skipped.
"""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SITE = os.path.join(ROOT, "janus_b200", "shims", "site")


def _ref_root():
    for cand in ("/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        if os.path.isdir(os.path.join(cand, "backend", "tests")):
            return cand
    pytest.skip("no reference tree (run scripts/stage_reference_tests.py where /root/reference is mounted)")


def _env(ref):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([SITE, ROOT, ref] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
    return env


def test_reference_test_suite_passes_over_the_shims(tmp_path):
    """backend/tests/test_input_processing.py (TestProsodyExtractor :436-505 runs REAL numerics through the YIN
    kernel; TestTranscriber :372-429; TestSenderMain :512-694), test_transport_layer.py, test_e2e_local.py,
    test_engine.py, test_api_flow.py, test_synthesis.py, test_voice_cloning.py -- as the reference ships them (58 tests)."""
    ref = _ref_root()
    tests = os.path.join(ref, "backend", "tests")
    # every test module the reference ships (hardware_check.py / manual_sender.py are manual scripts, not tests):
    # test_api_flow.py constructs the real `Transcriber()` in the FastAPI lifespan -- over the shim that is the B200 path
    files = [os.path.join(tests, f) for f in ("test_input_processing.py", "test_transport_layer.py", "test_e2e_local.py",
                                              "test_engine.py", "test_api_flow.py", "test_synthesis.py",
                                              "test_voice_cloning.py")]
    log = tmp_path / "maps.txt"
    # a tiny plugin: at session end, record which shared objects of this repo the pytest process has mapped
    plugin = tmp_path / "janus_maps_plugin.py"
    plugin.write_text(
        "def pytest_sessionfinish(session, exitstatus):\n"
        f"    open({str(log)!r}, 'w').write(''.join(l for l in open('/proc/self/maps') if 'libjanus_b200' in l))\n")
    env = _env(ref)
    env["PYTHONPATH"] = str(tmp_path) + os.pathsep + env["PYTHONPATH"]
    r = subprocess.run([sys.executable, "-m", "pytest", *files, "-q", "-p", "no:cacheprovider", "-p", "janus_maps_plugin"],
                       cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=900)
    tail = (r.stdout + r.stderr)[-3000:]
    print(tail)
    assert r.returncode == 0, tail
    m = re.search(r"(\d+) passed", r.stdout)
    assert m and int(m.group(1)) >= 55, tail
    assert "failed" not in r.stdout.splitlines()[-1]
    assert log.exists() and "libjanus_b200.so" in log.read_text(), "the reference's tests ran without the native library"


def test_reference_service_classes_unmodified_match_the_drop_in_and_the_oracle(tmp_path):
    """The reference's own `ProsodyExtractor` and `Transcriber` classes (their files, not ours) on the shims, in a
    subprocess: prosody tags equal the CPU oracle's on synthetic clips; the reference Transcriber's text equals the
    drop-in `janus_b200.services.transcriber.Transcriber` on the same 48 kHz buffer."""
    ref = _ref_root()
    code = r'''
import json, sys
import numpy as np
from backend.services.prosody import ProsodyExtractor as RefProsody          # the reference's file
from backend.services.transcriber import Transcriber as RefTranscriber       # the reference's file
import backend.services.prosody as P, backend.services.transcriber as T
assert "janus_b200" not in P.__file__ and "janus_b200" not in T.__file__
import aubio, faster_whisper
assert "shims" in aubio.__file__ and "shims" in faster_whisper.__file__
from janus_b200.synth import make_clips, sine_wave
from oracle.prosody_oracle import ProsodyOracle
out = {"prosody": [], "stt": {}}
clips = make_clips(3, 2.0, 48000, seed=61).numpy()
bufs = [clips[0], clips[1][:50001], clips[2], sine_wave(440.0, 0.5, 48000, 0.3), sine_wave(100.0, 0.5, 48000, 0.02)]
ref_p, orc = RefProsody(), ProsodyOracle()
for b in bufs:          # ONE stateful extractor on both sides (the aubio object carries its sliding buffer)
    out["prosody"].append([ref_p.analyze_buffer(b), orc.analyze_buffer(b, python_hop_loop=False)])
import janus_b200.services.transcriber as J
J.MAX_NEW_TOKENS = 32
rt = RefTranscriber("tiny")
rt.model._tr.max_windows = 1
ours = J.Transcriber("tiny"); ours.max_windows = 1
x48 = make_clips(1, 4.0, 48000, seed=63).numpy()[0]
chunks = [x48[i:i + 1536] for i in range(0, len(x48), 1536)]
out["stt"] = {"ref_class": rt.transcribe_buffer(chunks), "drop_in": ours.transcribe_buffer(x48)}
print("RESULT" + json.dumps(out))
'''
    r = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=_env(ref), capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    import json
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT")][-1][6:])
    for got, want in res["prosody"]:
        assert got == want, (got, want)
    assert {g["energy"] for g, _ in res["prosody"]} >= {"Quiet"}      # the cases cover more than one tag
    assert res["stt"]["ref_class"] == res["stt"]["drop_in"] and len(res["stt"]["drop_in"]) > 0
