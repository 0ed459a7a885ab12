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
    assert lib.janus_abi_version() == _abi.ABI_VERSION == 2
    assert lib.janus_yin_frames_per_group() >= 1


def test_every_declared_symbol_is_exported_and_bound():
    jbuild.build()
    raw = ctypes.CDLL(_abi.LIB_PATH)
    declared = _declared_symbols()
    assert declared, "no symbols parsed from include/*.h"
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in include/ but not exported"
    assert declared == set(_abi.SIGNATURES), (declared ^ set(_abi.SIGNATURES))


_SASS = {}


def _sass_of(pattern: str) -> dict:
    """function name -> list of SASS opcode mnemonics, for the kernels whose mangled name contains `pattern`."""
    import subprocess
    if "out" not in _SASS:
        _SASS["out"] = subprocess.run(["cuobjdump", "-sass", _abi.LIB_PATH], capture_output=True, text=True).stdout
    out = _SASS["out"]
    fns, cur = {}, None
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1) if pattern in m.group(1) else None
            if cur:
                fns[cur] = []
            continue
        if cur:
            m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
            if m:
                fns[cur].append(m.group(1))
    return fns


def test_sass_is_sm100a_and_yin_is_not_fma_contracted():
    """The bit-exact YIN claim rests on sub / mul / add being rounded separately, like aubio's scalar C loop compiled
    without contraction: the difference-function loop of yin_frames_kernel must hold FADD / FMUL and NO fused
    multiply-add (FFMA does appear elsewhere in the kernel, inside the correctly-rounded division sequences)."""
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    fns = _sass_of("yin_frames_kernel")
    assert fns, "yin_frames_kernel not found in the SASS"
    for name, ops in fns.items():
        # split the kernel at every FFMA: the unrolled difference-function loop must sit in ONE FFMA-free stretch that
        # holds >= 100 FMUL and two FADD per FMUL (x_j - x_{j+tau}, then the accumulate) -- i.e. nothing was contracted.
        # (The FFMAs that do exist come in groups of 5: the IEEE-correct division sequences of yin[tau] * tau / running_sum.)
        best, run = (0, 0), [0, 0]
        for o in ops + ["FFMA"]:
            if o.startswith("FFMA"):
                best = max(best, tuple(run))
                run = [0, 0]
            else:
                run[0] += o.startswith("FMUL")
                run[1] += o.startswith("FADD")
        fmul, fadd = best
        assert fmul >= 100 and abs(fadd - 2 * fmul) <= 8, (name, fmul, fadd)


def test_tensor_core_kernels_carry_tcgen05_and_tma_sass():
    """UTCHMMA (tcgen05.mma), LDTM/STTM (TMEM), UTMALDG (TMA) in the kernels DESIGN.md says use them."""
    for pat, need in (("gemm_bf16_kernel", ("UTCHMMA", "LDTM", "UTMALDG")),
                      ("attention_enc_kernel", ("UTCHMMA", "LDTM", "STTM", "UTMALDG")),
                      ("dec_xlat_kernel", ("UTCHMMA", "LDTM", "UTMALDG"))):
        fns = _sass_of(pat)
        assert fns, pat
        for name, ops in fns.items():
            for n in need:
                assert any(o.startswith(n) for o in ops), (name, n)


def test_invalid_args_fail_loudly_without_gpu():
    lib = _abi.lib()
    st = lib.janus_yin_frames_f32(1, 1, 1, 1, 1, 1, 1, 48000, 1000, 512, 0.8, -50.0, None, 1, 1, 1, None)
    assert st == -1 and b"buf_size" in lib.janus_last_error()
