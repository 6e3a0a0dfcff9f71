"""Host-side mirror of the reference's string functions, executed on the GPU.

Same names, argument meaning and error behaviour as the reference so callers (and
tests) read like the reference's own code:

    sanitize_query(query)              app.py:60-68
    is_safe_kubectl_command(command)   app.py:72-88
    KubectlOutputParser.parse(text)    app.py:95-104

plus the batched forms the single-string functions are built on.  Everything goes
through the C ABI (include/kx.h: kx_validate_hash / kx_validate_hash_host); there
is no Python implementation of the algorithms in this module.
"""
from __future__ import annotations

import logging
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib

logger = logging.getLogger(__name__)

SAFE, BAD_PREFIX, METACHAR, BAD_QUOTING = 0, 1, 2, 3
FLAG_PARSE = 1

# bytes handed to kx_validate_hash_host for the host -> device copy, and bytes it copies back, since import
# (counted here, at the call that performs the copies; bench.py reads the difference around a step)
copied = {"h2d": 0, "d2h": 0}


def to_csr(strings: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    """list[str] -> (uint8 data[sum len], int64 off[n+1]) in UTF-8.  Lone surrogates (which the reference's str
    functions accept but UTF-8 cannot carry) are encoded with 'surrogatepass': three bytes ED A0..BF xx, which the
    kernels treat as an ordinary 3-byte character, as str.split / strip / shlex do."""
    enc = [s.encode("utf-8", "surrogatepass") for s in strings]
    off = np.zeros(len(enc) + 1, np.int64)
    if enc:
        np.cumsum(np.fromiter((len(b) for b in enc), np.int64, len(enc)), out=off[1:])
    data = np.frombuffer(b"".join(enc), np.uint8)
    return data, off


def _ptr(a: Optional[np.ndarray]):
    return None if a is None or a.size == 0 else a.ctypes.data


def validate_hash_host(data: np.ndarray, off: np.ndarray, flags: int = 0, *, verdict=True, hash=True,
                       span=False, san=False, device: int = 0, out: Optional[dict] = None) -> dict:
    """kx_validate_hash_host over host CSR buffers.  Returns a dict of numpy arrays
    (verdict u8[n], hash i64[n], span i32[n,2], san u8[len(data)], san_len i32[n])
    for the outputs requested.  ``out`` may supply preallocated arrays."""
    data = np.ascontiguousarray(data, np.uint8)
    off = np.ascontiguousarray(off, np.int64)
    n = off.size - 1
    out = {} if out is None else out
    if verdict and "verdict" not in out:
        out["verdict"] = np.empty(n, np.uint8)
    if hash and "hash" not in out:
        out["hash"] = np.empty(n, np.int64)
    if span and "span" not in out:
        out["span"] = np.empty((n, 2), np.int32)
    if san and "san" not in out:
        out["san"] = np.empty(data.size, np.uint8)
        out["san_len"] = np.empty(n, np.int32)
    L = _lib.lib()
    copied["h2d"] += data.nbytes + off.nbytes
    copied["d2h"] += sum(out[k].nbytes for k in ("verdict", "hash", "span", "san", "san_len") if k in out)
    _lib.check(L.kx_validate_hash_host(_ptr(data), off.ctypes.data, n, flags, _ptr(out.get("verdict")),
                                       _ptr(out.get("hash")), _ptr(out.get("span")), _ptr(out.get("san")),
                                       _ptr(out.get("san_len")), device))
    return out


def validate_hash_device(data, off, flags: int = 0, *, verdict=None, hash=None, span=None, san=None,
                         san_len=None, stream=None) -> None:
    """kx_validate_hash over device-resident torch tensors (data u8, off i64[n+1]);
    outputs are written into the given tensors (None = skip).  Enqueues on
    ``stream`` (default: torch's current stream) and does not synchronise."""
    import torch

    n = off.numel() - 1
    st = torch.cuda.current_stream(data.device) if stream is None else stream

    def p(t):
        return None if t is None or t.numel() == 0 else t.data_ptr()

    with torch.cuda.device(data.device):
        _lib.check(_lib.lib().kx_validate_hash(p(data), off.data_ptr(), n, flags, p(verdict), p(hash), p(span),
                                               p(san), p(san_len), st.cuda_stream))


# --------------------------------------------------------------------------- #
# batched forms
# --------------------------------------------------------------------------- #
def command_verdicts(commands: Sequence[str], device: int = 0) -> np.ndarray:
    """Reason code (0 = safe) of is_safe_kubectl_command for every command."""
    data, off = to_csr(commands)
    return validate_hash_host(data, off, 0, hash=False, device=device)["verdict"]


def sanitize_queries(queries: Sequence[str], device: int = 0) -> Tuple[List[str], np.ndarray]:
    """(sanitize_query(q), hash(sanitize_query(q)) under PYTHONHASHSEED=0) per query."""
    data, off = to_csr(queries)
    r = validate_hash_host(data, off, 0, verdict=False, san=True, device=device)
    sb = r["san"].tobytes()
    lens = r["san_len"]
    return [sb[off[i]:off[i] + lens[i]].decode("utf-8", "surrogatepass") for i in range(len(queries))], r["hash"]


def parse_outputs(texts: Sequence[str], device: int = 0) -> Tuple[List[str], np.ndarray]:
    """What KubectlOutputParser.parse would validate, and its verdict, per text."""
    data, off = to_csr(texts)
    r = validate_hash_host(data, off, FLAG_PARSE, hash=False, span=True, device=device)
    raw = data.tobytes()
    lo = (off[:-1] + r["span"][:, 0]).tolist()
    hi = (off[:-1] + r["span"][:, 1]).tolist()
    cmds = [raw[a:b].decode("utf-8", "surrogatepass") for a, b in zip(lo, hi)]
    return cmds, r["verdict"]


# --------------------------------------------------------------------------- #
# single-string forms with the reference's names and behaviour
# --------------------------------------------------------------------------- #
def sanitize_query(query: str) -> str:
    """app.py:60-68."""
    return sanitize_queries([query])[0][0]


_REASON_LOG = {
    BAD_PREFIX: "Generated command does not start with 'kubectl ': %s",                 # app.py:76
    METACHAR: "Generated command contains potentially unsafe characters: %s",          # app.py:80
    BAD_QUOTING: "Generated command failed shlex parsing: %s",                          # app.py:86
}


def is_safe_kubectl_command(command: str) -> bool:
    """app.py:72-88 (logs a warning on rejection like the reference)."""
    v = int(command_verdicts([command])[0])
    if v != SAFE:
        logger.warning(_REASON_LOG[v], command.strip())
    return v == SAFE


class KubectlOutputParser:
    """Drop-in for the reference's parser (app.py:90-104): ``parse`` returns the
    command or raises ``ValueError`` with the reference's message."""

    def parse(self, text: str) -> str:
        cmds, verdicts = parse_outputs([text])
        if verdicts[0] != SAFE:
            raise ValueError(f"Generated command failed safety checks: {cmds[0]}")   # app.py:103
        return cmds[0]

    def parse_batch(self, texts: Sequence[str]) -> List[Optional[str]]:
        cmds, verdicts = parse_outputs(texts)
        return [c if v == SAFE else None for c, v in zip(cmds, verdicts)]
