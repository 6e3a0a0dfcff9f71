"""Deterministic byte-level tokenizer (no tokenizer vocabulary exists offline).

id 0..255 = the UTF-8 byte, 256 = BOS, 257 = EOS.  Random-init weights emit ids
anywhere in [0, vocab); ids above 257 decode to the byte ``id & 0xFF`` so that every
completion is a well-defined string for the parser/validator."""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

BOS, EOS = 256, 257


def encode(text: str, bos: bool = True) -> List[int]:
    ids = list(text.encode("utf-8"))
    return ([BOS] + ids) if bos else ids


def decode(ids: Sequence[int]) -> str:
    out = bytearray()
    for i in ids:
        i = int(i)
        if i == EOS:
            break
        if i == BOS:
            continue
        out.append(i & 0xFF)
    return out.decode("utf-8", errors="replace")


def encode_fixed(text: str, length: int, pad: int = 32) -> List[int]:
    """Prompt padded (left, with spaces) or truncated (keeping the tail, where the
    query and the 'Kubectl Command:' cue are) to exactly ``length`` tokens."""
    ids = encode(text, bos=False)
    if len(ids) >= length - 1:
        return [BOS] + ids[len(ids) - (length - 1):]
    return [BOS] + [pad] * (length - 1 - len(ids)) + ids


def encode_suffix(data: bytes, length: int, pad: int = 32) -> List[int]:
    """Bytes that continue a cached prefix, as exactly ``length`` tokens: left-padded with spaces or,
    if too long, truncated keeping the tail (no BOS: the prefix carries it)."""
    ids = list(data)
    if len(ids) >= length:
        return ids[len(ids) - length:]
    return [pad] * (length - len(ids)) + ids


# ---- batch versions used on the request path (same results as the scalar functions above) ----
def encode_fixed_batch(texts: Sequence[str], length: int, pad: int = 32) -> np.ndarray:
    out = np.full((len(texts), length), pad, np.int32)
    out[:, 0] = BOS
    for i, t in enumerate(texts):
        b = np.frombuffer(t.encode("utf-8"), np.uint8)
        if b.size > length - 1:
            b = b[b.size - (length - 1):]
        if b.size:
            out[i, length - b.size:] = b
    return out


def encode_suffix_batch(datas: Sequence[bytes], length: int, pad: int = 32) -> np.ndarray:
    out = np.full((len(datas), length), pad, np.int32)
    for i, d in enumerate(datas):
        b = np.frombuffer(d, np.uint8)
        if b.size > length:
            b = b[b.size - length:]
        if b.size:
            out[i, length - b.size:] = b
    return out


def decode_batch(ids: np.ndarray) -> List[str]:
    ids = np.asarray(ids)
    n = ids.shape[1]
    is_eos = ids == EOS
    end = np.where(is_eos.any(axis=1), is_eos.argmax(axis=1), n)
    bos = ids == BOS
    has_bos = (bos & (np.arange(n)[None, :] < end[:, None])).any(axis=1)
    low = (ids & 0xFF).astype(np.uint8)
    buf = low.tobytes()                      # one copy; rows without a BOS are sliced out of it (the usual case)
    out = []
    for i in range(ids.shape[0]):
        if has_bos[i]:
            out.append(low[i, :end[i]][~bos[i, :end[i]]].tobytes().decode("utf-8", errors="replace"))
        else:
            out.append(buf[i * n:i * n + end[i]].decode("utf-8", errors="replace"))
    return out
