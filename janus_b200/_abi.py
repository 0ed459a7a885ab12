"""ctypes binding of libjanus_b200.so (the C ABI declared in include/janus_b200.h).

No torch types cross this boundary: callers pass `tensor.data_ptr()` integers, explicit sizes and
the raw cudaStream_t handle. A non-zero status becomes a Python exception (the reference's callers
catch ordinary exceptions: /root/reference/backend/services/engine.py:513-525).
"""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libjanus_b200.so")

_c = ctypes
_P = _c.c_void_p
_I32 = _c.c_int32
_I64 = _c.c_int64
_F32 = _c.c_float

# name -> (restype, argtypes); must list every symbol include/janus_b200.h declares.
SIGNATURES = {
    "janus_abi_version": (_I32, []),
    "janus_last_error": (_c.c_char_p, []),
    "janus_check_device": (_I32, [_I32]),
    "janus_set_option": (_I32, [_I32, _I32]),
    "janus_yin_frames_per_group": (_I32, []),
    "janus_yin_frames_f32": (_I32, [_P, _P, _P, _P, _P, _I32, _I64, _I32, _I32, _I32, _F32, _F32,
                                    _P, _P, _P, _P, _P]),
    "janus_yin_state_f32": (_I32, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P]),
    "janus_prosody_reduce_f32": (_I32, [_P, _P, _P, _P, _I32, _P, _P, _P, _P, _P, _P]),
    "janus_prosody_reduce_static_f32": (_I32, [_P, _P, _P, _P, _I32, _I64, _P, _P, _P, _P, _P]),
    "janus_vad_gate_f32": (_I32, [_P, _P, _I64, _I32, _F32, _I32, _P, _P, _P]),
    "janus_layernorm_bf16": (_I32, [_P, _P, _P, _P, _I64, _I32, _P]),
    "janus_attention_enc_bf16": (_I32, [_P, _P, _I32, _I32, _I32, _P]),
    "janus_logmel_frames_alloc": (_I32, []),
    "janus_logmel_f32": (_I32, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P, _I32, _P, _P, _P, _P, _P]),
    "janus_logmel_utterance_f32": (_I32, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P, _I32, _P, _I32, _P, _P]),
    "janus_logmel_windows_f32": (_I32, [_P, _P, _P, _I32, _I32, _I32, _P, _P, _I32, _P, _P, _P]),
    "janus_whisper_encode_workspace_bytes": (_I64, [_P, _I32]),
    "janus_whisper_encode_bf16": (_I32, [_P, _P, _P, _I32, _P, _P, _I64, _P]),
    "janus_dec_linear_bf16": (_I32, [_P, _I32, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _P]),
    "janus_dec_cross_attn_bf16": (_I32, [_P, _P, _P, _P, _I32, _I32, _I32, _I32, _P]),
    "janus_dec_xlat_bf16": (_I32, [_P, _P, _P, _I32, _I32, _I32, _I32, _P]),
    "janus_whisper_cross_kv_bf16": (_I32, [_P, _P, _P, _I32, _P, _P, _P]),
    "janus_whisper_decode_step": (_I32, [_P, _P, _P, _P]),
    "janus_whisper_decode_beam_step": (_I32, [_P, _P, _P, _P, _P]),
    "janus_whisper_beam_finalize": (_I32, [_P, _P, _I32, _P, _P, _P, _P]),
    "janus_dec_cross_attn_beam_bf16": (_I32, [_P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _I32, _P]),
    "janus_whisper_cross_kv_swapped_bf16": (_I32, [_P, _P, _P, _I32, _P, _P, _P]),
    "janus_dec_xlat_grouped_bf16": (_I32, [_P, _P, _P, _I32, _I32, _I32, _I32, _I32, _P]),
    "janus_gemm_bf16": (_I32, [_P, _I64, _I64, _P, _I64, _P, _I64, _I64, _I32, _I32, _I32, _I32, _I32,
                               _P, _P, _P, _I32, _P]),
}


ABI_VERSION = 2   # JANUS_ABI_VERSION of include/janus_b200.h these bindings were written against

# janus_option_t (include/janus_b200.h)
OPT_PDL, OPT_ATT_SPLIT, OPT_XLAT_STREAM, OPT_XLAT_CTAS, OPT_XA_CTAS_PER_SM, OPT_BEAM_SHARED_SELF_ATTN = range(6)


def set_option(option: int, value: int) -> None:
    """Process-wide tunable of the library (value < 0 restores the default); see janus_option_t."""
    call("janus_set_option", option, value)


class JanusNativeError(RuntimeError):
    """Raised when a libjanus_b200 entry point returns a non-zero status."""


_lib = None
_lock = threading.Lock()


def lib() -> ctypes.CDLL:
    """Loads libjanus_b200.so, building it with nvcc first if it is missing or stale.

    There is no fallback: if the library cannot be built or loaded this raises.
    """
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        from . import build as _build
        try:
            _build.build()       # a no-op unless the library is missing or older than its sources
        except Exception as exc:  # noqa: BLE001
            if not os.path.exists(LIB_PATH):
                raise JanusNativeError(
                    f"libjanus_b200.so is missing and could not be built with nvcc: {exc}. "
                    "Run `python -m janus_b200.build`. There is no CPU fallback.") from exc
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        if L.janus_abi_version() != ABI_VERSION:
            raise JanusNativeError(f"libjanus_b200.so ABI version {L.janus_abi_version()} != {ABI_VERSION} expected by "
                                   "janus_b200/_abi.py: rebuild with `python -m janus_b200.build --force`")
        _lib = L
    return _lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().janus_last_error()
        raise JanusNativeError(f"{what} failed (status {status}): {msg.decode() if msg else ''}")


def call(name: str, *args) -> None:
    check(getattr(lib(), name)(*args), name)
