"""ctypes binding of libkx_b200.so (the C ABI declared in include/kx.h).

There is no CPU fallback: if the CUDA library cannot be loaded this raises."""
from __future__ import annotations

import ctypes
import os

from . import build as _build

_LIB = None

c_void_p, c_int, c_int64, c_uint32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint32


class KxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libkx_b200 error {code}: {msg}")
        self.code = code


_SIGNATURES = {
    "kx_last_error": (ctypes.c_char_p, []),
    "kx_version": (c_int, []),
    "kx_launch_count": (c_int64, []),
    "kx_validate_hash": (c_int, [c_void_p, c_void_p, c_int64, c_uint32] + [c_void_p] * 6),
    "kx_validate_hash_noalloc": (c_int, [c_void_p, c_void_p, c_int64, c_uint32] + [c_void_p] * 6),
    "kx_validate_hash_host": (c_int, [c_void_p, c_void_p, c_int64, c_uint32] + [c_void_p] * 5 + [c_int]),
    "kx_release_host_staging": (None, []),
    "kx_release_scratch": (None, []),
    "kx_gemm_bf16": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int, c_int, c_int, c_void_p,
                             c_int64, c_void_p, c_int, c_void_p]),
    "kx_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, ctypes.c_float, c_void_p, c_void_p, c_int, c_int64, c_void_p]),
    "kx_argmax_parts": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "kx_attn_decode": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]
                       + [c_int] * 7 + [c_void_p]),
}


def declare(name, restype, argtypes):
    """Later modules (generator) register their own entry points here."""
    _SIGNATURES[name] = (restype, argtypes)
    if _LIB is not None:
        fn = getattr(_LIB, name)
        fn.restype, fn.argtypes = restype, argtypes


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        path = _build.LIB_PATH
        if not os.path.exists(path):
            try:
                _build.build_lib()
            except Exception as e:  # noqa: BLE001
                raise ImportError(
                    f"{path} is missing and could not be built ({e}); run `python __graft_entry__.py build`. "
                    "This package has no CPU fallback.") from e
        L = ctypes.CDLL(path)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(L, name)   # AttributeError here means header and library disagree
            fn.restype, fn.argtypes = restype, argtypes
        _LIB = L
    return _LIB


def check(rc: int) -> None:
    if rc != 0:
        raise KxError(rc, lib().kx_last_error().decode("utf-8", "replace"))
