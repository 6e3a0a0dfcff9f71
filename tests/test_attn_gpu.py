"""Decode attention kernels through kx_attn_decode against a plain PyTorch fp32
reference of the same op: RoPE of the new token's q / k, append to the paged cache, softmax(q K^T) V
over a ragged batch.  The shapes steer the library into each of its kernels: one CTA per
(sequence, kv head) with the keys split over its warps, one warp per pair, and the shared-prefix
path (prefix keys reduced once per 64 query rows, merged with each sequence's own keys)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
PAGE = 16


def _rope_table(max_pos, hd, theta=10000.0):
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float32) / hd))
    ang = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv[None, :]
    return torch.stack([ang.cos(), ang.sin()], dim=-1).contiguous()          # [max_pos, hd/2, 2]


def _rot(x, cs):                                                              # x [..., hd] fp32, cs [hd/2, 2]
    half = x.shape[-1] // 2
    x1, x2 = x[..., :half], x[..., half:]
    return torch.cat([x1 * cs[:, 0] - x2 * cs[:, 1], x2 * cs[:, 0] + x1 * cs[:, 1]], dim=-1)


def _case(B, n_q, n_kv, hd, lens, prefix_len, seed):
    """lens[b] = keys already cached for sequence b (the new token lands at position lens[b])."""
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import _lib
    g = torch.Generator().manual_seed(seed)
    G = n_q // n_kv
    max_T = max(lens) + 1
    max_pages = (max_T + PAGE - 1) // PAGE
    rows = B + 1                                                               # last row: the shared prefix
    num_pages = rows * max_pages
    pages = (torch.randn(num_pages, 2, n_kv, PAGE, hd, generator=g) * 0.7).to(torch.bfloat16)
    perm = torch.randperm(num_pages, generator=g).to(torch.int32)              # pages scattered over the pool
    bt = perm.view(rows, max_pages).clone()
    if prefix_len:
        bt[:B, :prefix_len // PAGE] = bt[B, :prefix_len // PAGE]
    qkv = (torch.randn(B, (n_q + 2 * n_kv) * hd, generator=g) * 0.8).to(torch.bfloat16)
    rope = _rope_table(max_T + 8, hd)
    pos = torch.tensor(lens, dtype=torch.int32)

    # ---- fp32 reference ----
    want = torch.zeros(B, n_q, hd)
    qv = qkv.float().view(B, n_q + 2 * n_kv, hd)
    k_new = torch.zeros(B, n_kv, hd)
    for b in range(B):
        p = lens[b]
        qr = _rot(qv[b, :n_q], rope[p]).to(torch.bfloat16).float()
        kr = _rot(qv[b, n_q:n_q + n_kv], rope[p]).to(torch.bfloat16).float()
        k_new[b] = kr
        vr = qv[b, n_q + n_kv:]
        idx = bt[b, torch.arange(p) // PAGE].long()
        K = pages[idx, 0, :, torch.arange(p) % PAGE].float()                  # [p, n_kv, hd]
        V = pages[idx, 1, :, torch.arange(p) % PAGE].float()
        K = torch.cat([K, kr[None]], 0)
        V = torch.cat([V, vr[None]], 0)
        for h in range(n_q):
            s = (K[:, h // G] @ qr[h]) / math.sqrt(hd)
            want[b, h] = torch.softmax(s, 0) @ V[:, h // G]

    d_pages, d_bt, d_qkv, d_rope, d_pos = (t.cuda() for t in (pages, bt, qkv, rope, pos))
    out = torch.zeros(B, n_q * hd, dtype=torch.bfloat16, device="cuda")
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):                                                         # second call: cache already holds the new token
        _lib.check(L.kx_attn_decode(d_qkv.data_ptr(), d_qkv.stride(0), out.data_ptr(), out.stride(0), d_pages.data_ptr(),
                                    d_bt.data_ptr(), max_pages, d_pos.data_ptr(), d_rope.data_ptr(), B, max_T, n_q, n_kv, hd,
                                    prefix_len, B, st))
        torch.cuda.synchronize()
        got = out.float().cpu().view(B, n_q, hd)
        err = (got - want).abs().max().item() / want.abs().max().item()
        assert err <= 1e-2, f"attention output differs from the fp32 reference by {err:.2e} of its scale"
    # the new token's k (rotated) and v were appended to the cache
    hp = d_pages.cpu()
    for b in range(B):
        p = lens[b]
        page = bt[b, p // PAGE].item()
        assert torch.equal(hp[page, 0, :, p % PAGE].float(), k_new[b]), f"appended k of sequence {b}"
        assert torch.equal(hp[page, 1, :, p % PAGE], qkv[b].view(n_q + 2 * n_kv, hd)[n_q + n_kv:]), f"appended v of sequence {b}"
    return err


def _ragged(B, lo, hi, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(lo, hi + 1, (B,), generator=g).tolist()


@pytest.mark.parametrize("hd,n_q,n_kv", [(64, 8, 2), (128, 4, 1), (64, 14, 2), (64, 2, 2)])
def test_cta_per_pair_kernel_long_ragged_context(hd, n_q, n_kv):
    _case(5, n_q, n_kv, hd, _ragged(5, 130, 900, 1), 0, 1)                     # few pairs, many keys: CTA kernel, several rounds


@pytest.mark.parametrize("hd,n_q,n_kv", [(64, 8, 2), (128, 4, 1), (64, 14, 2), (128, 8, 8)])
def test_warp_per_pair_kernel_short_context(hd, n_q, n_kv):
    _case(9, n_q, n_kv, hd, _ragged(9, 0, 120, 2), 0, 2)                       # <= 128 keys: one warp per pair (deep ring)


@pytest.mark.parametrize("hd,n_q,n_kv", [(64, 32, 8), (128, 8, 4)])
def test_warp_per_pair_kernel_wide_batch(hd, n_q, n_kv):
    B = 512 // n_kv + 3
    _case(B, n_q, n_kv, hd, _ragged(B, 200, 330, 3), 0, 3)                     # >= 512 pairs: one warp per pair, 2-stage ring


@pytest.mark.parametrize("hd,n_q,n_kv,B,P,tail", [(128, 16, 4, 64, 1024, 96), (64, 8, 2, 40, 336, 40), (64, 14, 2, 3, 64, 10),
                                                    (128, 2, 1, 70, 208, 300), (64, 32, 8, 130, 80, 30)])
def test_shared_prefix_path(hd, n_q, n_kv, B, P, tail):
    lens = [P + t for t in _ragged(B, 0, tail, 4)]                              # every sequence: the prefix + its own keys
    _case(B, n_q, n_kv, hd, lens, P, 4)
