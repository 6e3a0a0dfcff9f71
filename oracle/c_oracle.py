"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/kx_oracle.c (built by
``make -C oracle`` or ``__graft_entry__.build()``)."""
from __future__ import annotations

import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libkx_oracle.so")
    src = os.path.join(_HERE, "kx_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libkx_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.kx_oracle_validate_hash.restype = ctypes.c_int
        _LIB.kx_oracle_validate_hash.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                 ctypes.c_int64, ctypes.c_uint32] + [ctypes.c_void_p] * 5
    return _LIB


def validate_hash(data: np.ndarray, off: np.ndarray, flags: int = 0, want_span=False, want_san=False,
                  threads: int = 1):
    """Returns dict(verdict u8[n], hash i64[n], span i32[n,2]?, san u8[len(data)]?, san_len i32[n]?)."""
    data = np.ascontiguousarray(data, dtype=np.uint8)
    off = np.ascontiguousarray(off, dtype=np.int64)
    n = off.size - 1
    out = {"verdict": np.empty(n, np.uint8), "hash": np.empty(n, np.int64)}
    if want_span:
        out["span"] = np.empty((n, 2), np.int32)
    if want_san:
        out["san"] = np.zeros(max(1, data.size), np.uint8)
        out["san_len"] = np.empty(n, np.int32)
    L = lib()

    def ptr(a):
        return a.ctypes.data if a is not None else None

    def run(r):
        rc = L.kx_oracle_validate_hash(ptr(data) if data.size else None, ptr(off), r[0], r[1], flags,
                                       ptr(out["verdict"]), ptr(out["hash"]), ptr(out.get("span")),
                                       ptr(out.get("san")), ptr(out.get("san_len")))
        if rc:
            raise MemoryError("kx_oracle_validate_hash")

    threads = max(1, min(threads, n or 1))
    cuts = [n * t // threads for t in range(threads + 1)]
    ranges = [(cuts[t], cuts[t + 1]) for t in range(threads)]
    if threads == 1:
        run(ranges[0])
    else:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(run, ranges))
    return out
