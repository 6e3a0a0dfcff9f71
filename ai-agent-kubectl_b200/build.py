"""Builds libkx_b200.so in-tree with nvcc for sm_100a (no JIT cache, so the .so
travels with the repo snapshot to the GPU box).  Every .cu file is compiled to its own
object (in parallel, cached by modification time under .obj/) and the objects are linked."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
OBJ_DIR = os.path.join(PKG_DIR, ".obj")
LIB_PATH = os.path.join(PKG_DIR, "libkx_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: libkx_b200.so cannot be built")
    return p


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def headers():
    return glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(os.path.dirname(PKG_DIR), "include", "*.h"))


def stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in sources() + headers())


def _compile(src: str, force: bool, extra, verbose: bool) -> str:
    obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
    newest = max(os.path.getmtime(p) for p in [src] + headers())
    if not force and not extra and os.path.exists(obj) and os.path.getmtime(obj) > newest:
        return obj
    cmd = [nvcc_path()] + NVCC_FLAGS + list(extra) + ["-c", "-o", obj, src]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return obj


def build_lib(force: bool = False, verbose: bool = False, extra=()) -> str:
    if not force and not stale():
        return LIB_PATH
    # several ranks of one job may get here at once (torchrun): one of them builds, into a temporary file that
    # replaces the library atomically; the others wait on the lock and then find it up to date
    import fcntl
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if force or stale():
            os.makedirs(OBJ_DIR, exist_ok=True)
            with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
                objs = list(pool.map(lambda s: _compile(s, force, extra, verbose), sources()))
            tmp = LIB_PATH + f".tmp{os.getpid()}"
            cmd = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", tmp] + objs + ["-lcuda"]
            if verbose:
                print(" ".join(cmd), flush=True)
            try:
                subprocess.check_call(cmd)
                os.replace(tmp, LIB_PATH)
            finally:
                if os.path.exists(tmp):
                    os.remove(tmp)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    build_lib(force="--force" in sys.argv, verbose=True, extra=[a for a in sys.argv[1:] if a != "--force"])
