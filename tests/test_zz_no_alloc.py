"""The allocation-free entry point (kx_validate_hash_noalloc) against the regular one.

Some hosts of this library cannot allocate device memory at call time: all of HBM may
already belong to a serving framework's allocator, or the request loop cannot tolerate
the device-wide synchronisation a cudaMalloc causes.  For them the library offers a
variant of kx_validate_hash whose scratch queue is a fixed array inside the library.
This file checks, through the C ABI, that it returns exactly what the regular entry
point returns -- on two calls in a row, because the array is shared and keeps the
previous call's contents.  It is named so that pytest collects it last.
"""
import ctypes

import numpy as np
import pytest


def _lib():
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import _lib
    return _lib


def test_batches_beyond_the_fixed_capacity_are_rejected():
    L = _lib().lib()
    too_many = (1 << 20) + 1
    rc = L.kx_validate_hash_noalloc(None, None, too_many, 0, None, None, None, None, None, None)
    assert rc == -4                                      # KX_ETOOBIG
    msg = L.kx_last_error().decode()
    assert "1048576" in msg and "kx_validate_hash" in msg


@pytest.mark.gpu
def test_two_calls_in_a_row_match_the_allocating_entry_point():
    import torch
    from kubectl_b200 import synth
    lib = _lib()
    L = lib.lib()
    stream = torch.cuda.current_stream().cuda_stream

    def run(fn, d_data, d_off, n, flags):
        out = {"verdict": torch.full((n,), 255, dtype=torch.uint8, device="cuda"),
               "hash": torch.zeros(n, dtype=torch.int64, device="cuda"),
               "span": torch.full((2 * n,), -1, dtype=torch.int32, device="cuda"),
               "san": torch.zeros(max(1, d_data.numel()), dtype=torch.uint8, device="cuda"),
               "san_len": torch.full((n,), -1, dtype=torch.int32, device="cuda")}
        lib.check(fn(d_data.data_ptr(), d_off.data_ptr(), n, flags, out["verdict"].data_ptr(),
                     out["hash"].data_ptr(), out["span"].data_ptr(), out["san"].data_ptr(),
                     out["san_len"].data_ptr(), stream))
        torch.cuda.synchronize()
        return {k: v.cpu().numpy() for k, v in out.items()}

    # call 1: every string is non-ASCII, so the whole batch goes through the scratch queue;
    # call 2: a smaller, mostly-ASCII batch in PARSE mode -- stale queue contents from call 1
    # would show up as wrong (or out-of-range) work if the counter were not cleared.
    first = [s + " é€\U0001f600" for s in synth.command_strings(30_000, seed=11)]
    second = synth.command_strings(7_000, seed=12) + ["```\nkubectl get pods\n```", "  kubectl get ns  "]
    for strings, flags in ((first, 0), (second, 1)):
        enc = [s.encode() for s in strings]
        off = np.zeros(len(enc) + 1, np.int64)
        off[1:] = np.cumsum([len(b) for b in enc])
        d_data = torch.from_numpy(np.frombuffer(b"".join(enc), np.uint8).copy()).cuda()
        d_off = torch.from_numpy(off).cuda()
        n = len(enc)
        want = run(L.kx_validate_hash, d_data, d_off, n, flags)
        got = run(L.kx_validate_hash_noalloc, d_data, d_off, n, flags)
        for k in ("verdict", "hash", "span", "san_len"):          # integer work: bit-exact
            assert np.array_equal(got[k], want[k]), k
        assert (want["verdict"] != 255).all()
        for i in range(n):
            a, ln = off[i], want["san_len"][i]
            assert np.array_equal(got["san"][a:a + ln], want["san"][a:a + ln])
