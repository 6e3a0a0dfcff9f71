"""Builds libkx_b200.so in-tree with nvcc for sm_100a (no JIT cache, so the .so
travels with the repo snapshot to the GPU box)."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libkx_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr"]


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: libkx_b200.so cannot be built")
    return p


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(os.path.dirname(PKG_DIR), "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build_lib(force: bool = False, verbose: bool = False, extra=()) -> str:
    if not force and not stale():
        return LIB_PATH
    # several ranks of one job may get here at once (torchrun): one of them builds, into a temporary file that
    # replaces the library atomically; the others wait on the lock and then find it up to date
    import fcntl
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if force or stale():
            tmp = LIB_PATH + f".tmp{os.getpid()}"
            cmd = [nvcc_path()] + NVCC_FLAGS + list(extra) + ["-o", tmp] + sources() + ["-lcuda"]
            if verbose:
                print(" ".join(cmd))
            try:
                subprocess.check_call(cmd)
                os.replace(tmp, LIB_PATH)
            finally:
                if os.path.exists(tmp):
                    os.remove(tmp)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    build_lib(force=True, verbose=True, extra=[a for a in sys.argv[1:]])
