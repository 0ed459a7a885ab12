"""The C-ABI library builds, loads, and exports every symbol include/janus_b200.h declares.
No compute calls here (no GPU needed)."""
import ctypes
import os
import re

from janus_b200 import _abi
from janus_b200 import build as jbuild

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(janus_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_builds_and_loads():
    path = jbuild.build()
    assert os.path.exists(path)
    lib = _abi.lib()
    assert lib.janus_abi_version() == 1
    assert lib.janus_yin_frames_per_group() >= 1


def test_every_declared_symbol_is_exported_and_bound():
    jbuild.build()
    raw = ctypes.CDLL(_abi.LIB_PATH)
    declared = _declared_symbols()
    assert declared, "no symbols parsed from include/*.h"
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in include/ but not exported"
    assert declared == set(_abi.SIGNATURES), (declared ^ set(_abi.SIGNATURES))


def test_sass_is_sm100a_and_yin_is_not_fma_contracted():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_invalid_args_fail_loudly_without_gpu():
    lib = _abi.lib()
    st = lib.janus_yin_frames_f32(1, 1, 1, 1, 1, 1, 1, 48000, 1000, 512, 0.8, -50.0, None, 1, 1, 1, None)
    assert st == -1 and b"buf_size" in lib.janus_last_error()
