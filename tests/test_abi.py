"""CPU-side checks of the C ABI: the library builds for sm_100a, loads, and exports
every symbol include/kx.h declares.  No compute calls (there is no GPU here)."""
import ctypes
import os
import re

from conftest import ROOT

import kubectl_b200  # noqa: F401
from kubectl_b200 import _lib, build, engine  # noqa: F401  (engine registers its entry points)


def declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            text = open(os.path.join(ROOT, "include", fn)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            names |= set(re.findall(r"\b(kx_\w+)\s*\(", text))
    return names


def test_library_builds_and_exports_every_declared_symbol():
    path = build.build_lib()
    assert os.path.exists(path)
    L = ctypes.CDLL(path)
    syms = declared_symbols()
    assert {"kx_validate_hash", "kx_validate_hash_host", "kx_last_error"} <= syms
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/ but not exported"


def test_binding_declares_every_symbol():
    _lib.lib()
    missing = declared_symbols() - set(_lib._SIGNATURES)
    assert not missing, f"ctypes binding lacks {missing}"


def test_argument_errors_are_reported_without_a_gpu():
    L = _lib.lib()
    assert L.kx_validate_hash(None, None, -1, 0, None, None, None, None, None, None) == -1
    assert b"negative" in L.kx_last_error()
    assert L.kx_validate_hash(None, None, 1, 0x80, None, None, None, None, None, None) == -1
    assert b"flags" in L.kx_last_error()
    assert L.kx_validate_hash(None, None, 0, 0, None, None, None, None, None, None) == 0


def test_sass_is_sm100a():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", build.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
