"""Runs the CUDA kernel's per-string logic (k9_core.cuh, host+device code) on the
CPU against the oracle.  This is a check of the kernel source, not a product path:
the shipped library only ever executes it on the GPU."""
import ctypes
import os
import random
import subprocess

import numpy as np
import pytest

from conftest import ROOT, to_csr
from oracle import c_oracle


@pytest.fixture(scope="module")
def emul():
    src = os.path.join(ROOT, "tests", "emul", "k9_emul.cpp")
    so = os.path.join(ROOT, "tests", "emul", "k9_emul.so")
    core = os.path.join(ROOT, "ai-agent-kubectl_b200", "csrc", "k9_core.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(core)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", src, "-o", so])
    L = ctypes.CDLL(so)
    L.k9_emul.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_uint32] + [ctypes.c_void_p] * 5 + [ctypes.c_int]
    L.k9_emul.restype = ctypes.c_int64

    def run(data, off, flags, mode=0):
        n = off.size - 1
        out = {"verdict": np.empty(n, np.uint8), "hash": np.empty(n, np.int64), "span": np.empty((n, 2), np.int32),
               "san": np.zeros(max(1, data.size), np.uint8), "san_len": np.empty(n, np.int32)}
        data = np.ascontiguousarray(data)
        out["nslow"] = L.k9_emul(data.ctypes.data if data.size else None, off.ctypes.data, n, flags,
                                 *(out[k].ctypes.data for k in ("verdict", "hash", "span", "san", "san_len")), mode)
        return out
    return run


def _same(a, b, off):
    for k in ("verdict", "hash", "span", "san_len"):
        bad = np.flatnonzero((a[k] != b[k]).reshape(len(a[k]), -1).any(axis=1))
        assert bad.size == 0, (k, bad[:5], a[k][bad[:5]], b[k][bad[:5]])
    for i in np.flatnonzero(a["san_len"] > 0):
        s, n = off[i], a["san_len"][i]
        assert np.array_equal(a["san"][s:s + n], b["san"][s:s + n]), i


@pytest.mark.parametrize("flags", [0, 1])
def test_kernel_logic_on_golden_strings(emul, golden_rows, flags):
    data, off = to_csr([r["s"] for r in golden_rows])
    got = emul(data, off, flags)
    _same(got, c_oracle.validate_hash(data, off, flags, True, True), off)
    if flags == 0:
        assert [int(v) for v in got["verdict"]] == [r["reason"] for r in golden_rows]
        assert [int(h) for h in got["hash"]] == [r["key_hash"] for r in golden_rows]


@pytest.mark.parametrize("flags", [0, 1])
def test_kernel_logic_on_synthetic_and_garbage(emul, flags):
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import synth
    data, off = synth.command_csr(50_000, seed=2)
    ref = c_oracle.validate_hash(data, off, flags, True, True)
    got = emul(data, off, flags)
    _same(got, ref, off)
    assert 0 < got["nslow"] < 0.1 * 50_000          # only the non-ASCII strings leave the fast path
    _same(emul(data, off, flags, mode=1), ref, off)  # general path alone
    rng = random.Random(5)
    al = [0x20, 0x27, 0x22, 0x5C, 0x60, 0x6B, 0x80, 0xA0, 0xC2, 0xE2, 0xE3, 0xF0, 0xFF, 0x28, 0x85, 0x9F, 0x81, 0x00,
          0x40, 0xA8, 0xC1, 0x09]
    blobs = [b"kubectl " * (rng.random() < 0.5) + bytes(rng.choice(al) for _ in range(rng.randrange(0, 40)))
             for _ in range(30000)]
    off = np.zeros(len(blobs) + 1, np.int64)
    off[1:] = np.cumsum([len(b) for b in blobs])
    data = np.frombuffer(b"".join(blobs), np.uint8)
    _same(emul(data, off, flags), c_oracle.validate_hash(data, off, flags, True, True), off)


@pytest.mark.parametrize("flags", [0, 1])
def test_unicode_whitespace_at_the_ends_stays_on_the_fast_path(emul, flags):
    """str.strip() removes non-ASCII whitespace too (app.py:74,97).  The fast path strips it at both ends when
    the rest is ASCII; pieces of broken UTF-8 next to the padding must still give exactly the oracle's answer
    (they change where the general path's forward decoder finds code point boundaries)."""
    rng = random.Random(11)
    ws = [" ", "\t", "\x85", "\xa0", "\u1680", "\u2000", "\u2005", "\u200a", "\u2028", "\u2029", "\u202f", "\u205f", "\u3000"]
    ws = [w.encode("utf-8") for w in ws]
    junk = [b"\xe2\x80", b"\x80", b"\xc2", b"\xf0\xe2\x80\xa8", b"\xe2", b"\xe3\x80", b"\xc2\x84", b"\xe2\x80\x8b", b"\xe1\x9a"]
    cores = [b"kubectl get pods", b"kubectl get pods -l 'a b'", b"kubectl", b"", b"kubectl get x; y", b"```\nkubectl get ns\n```",
             b"kubectl get pods \"unclosed", b"k"]

    def pad():
        out = b""
        for _ in range(rng.randrange(0, 4)):
            out += rng.choice(ws) if rng.random() < 0.8 else rng.choice(junk)
        return out
    blobs = [pad() + rng.choice(cores) + pad() for _ in range(40000)]
    off = np.zeros(len(blobs) + 1, np.int64)
    off[1:] = np.cumsum([len(b) for b in blobs])
    data = np.frombuffer(b"".join(blobs), np.uint8)
    got = emul(data, off, flags)
    _same(got, c_oracle.validate_hash(data, off, flags, True, True), off)
    clean = sum(1 for b in blobs if not any(j in b for j in junk))
    assert got["nslow"] <= len(blobs) - 0.9 * clean + (0.2 * len(blobs) if flags else 0)   # padded ASCII strings did not leave it
