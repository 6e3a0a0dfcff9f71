"""tcgen05 GEMM (k1/k3) against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gemm():
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import _lib

    def run(a, b, epi, swap, out, bias=None, splits=1):
        L = _lib.lib()
        _lib.check(L.kx_gemm_bf16(a.data_ptr(), a.shape[0], a.stride(0), b.data_ptr(), b.shape[0], b.stride(0),
                                  a.shape[1], epi, swap, out.data_ptr(), out.stride(0),
                                  None if bias is None else bias.data_ptr(), splits,
                                  torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
    return run


def _rand(r, c, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(r, c, generator=g, device="cuda") * 0.5).to(torch.bfloat16)


def _ref(a, b):
    return a.float() @ b.float().T


SHAPES = [(128, 128, 64), (300, 256, 192), (1000, 1152, 896), (257, 144, 4864), (4096, 2048, 896), (2049, 896, 512),
          (640, 9728, 896),
          (520, 4352, 8192)]      # 17 feature tiles of 4 MB: three bands of the pair kernel's tile walk, the last one narrower


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_prefill_orientation_all_epilogues(gemm, m, n, k):
    a, w = _rand(m, k, 1), _rand(n, k, 2)
    bias = torch.randn(n, device="cuda")
    ref = _ref(a, w)
    out = torch.zeros(m, n, device="cuda", dtype=torch.bfloat16)
    gemm(a, w, 0, 0, out, bias)
    torch.testing.assert_close(out.float(), (ref + bias).to(torch.bfloat16).float(), rtol=2e-2, atol=2e-2)
    f32 = torch.zeros(m, n, device="cuda")
    gemm(a, w, 3, 0, f32)
    torch.testing.assert_close(f32, ref, rtol=1e-4, atol=1e-3)
    resid = torch.randn(m, n, device="cuda")
    want = resid + ref
    gemm(a, w, 1, 0, resid)
    torch.testing.assert_close(resid, want, rtol=1e-4, atol=1e-3)
    sw = torch.zeros(m, n // 2, device="cuda", dtype=torch.bfloat16)
    gemm(a, w, 2, 0, sw)
    want = torch.nn.functional.silu(ref[:, 0::2]) * ref[:, 1::2]
    torch.testing.assert_close(sw.float(), want, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("batch", [1, 20, 64, 100, 200])
@pytest.mark.parametrize("feat,k,splits", [(384, 896, 1), (1152, 896, 7), (896, 4864, 19), (1024, 128, 4)])
def test_decode_orientation_swap_ab_and_split_k(gemm, batch, feat, k, splits):
    w, x = _rand(feat, k, 3), _rand(batch, k, 4)
    bias = torch.randn(feat, device="cuda")
    ref = _ref(x, w)                                    # [batch, feat]
    for rep in range(2):                                # scratch must come back zeroed
        out = torch.zeros(batch, feat, device="cuda", dtype=torch.bfloat16)
        gemm(w, x, 0, 1, out, bias, splits)
        torch.testing.assert_close(out.float(), (ref + bias).to(torch.bfloat16).float(), rtol=2e-2, atol=2e-2)
        f32 = torch.zeros(batch, feat, device="cuda")
        gemm(w, x, 3, 1, f32, None, splits)
        torch.testing.assert_close(f32, ref, rtol=1e-4, atol=2e-3)
        resid = torch.randn(batch, feat, device="cuda")
        want = resid + ref
        gemm(w, x, 1, 1, resid, None, splits)
        torch.testing.assert_close(resid, want, rtol=1e-4, atol=2e-3)
        sw = torch.zeros(batch, feat // 2, device="cuda", dtype=torch.bfloat16)
        gemm(w, x, 2, 1, sw, None, splits)
        want = torch.nn.functional.silu(ref[:, 0::2]) * ref[:, 1::2]
        torch.testing.assert_close(sw.float(), want, rtol=2e-2, atol=2e-2)


def test_strided_operands(gemm):
    big = _rand(300, 1024, 5)
    a = big[:, 128:128 + 256]                            # row pitch 1024, k = 256
    w = _rand(128, 256, 6)
    out = torch.zeros(300, 128, device="cuda")
    gemm(a, w, 3, 0, out)
    torch.testing.assert_close(out, _ref(a, w), rtol=1e-4, atol=1e-3)


def _argmax_rows(ref):
    """(max, lowest index) per row on the CPU: the tie rule of torch.argmax / the reference decoder."""
    import numpy as np
    a = ref.cpu().numpy()
    return np.argmax(a, axis=1)          # numpy returns the first maximal index


def _check_against_torch(ids, x, w):
    """... and equal to torch.argmax over a plain PyTorch fp32 product of the same bf16 operands wherever that product's
    top-2 margin is clear of fp32 summation-order noise (the exact-tie row is covered by the comparison above)."""
    ref = _ref(x, w)
    top2 = ref.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 1e-4 * ref.abs().max()
    assert clear.float().mean() > 0.9
    assert torch.equal(ids.long()[clear], ref.argmax(dim=1)[clear])


@pytest.mark.parametrize("batch,vocab,k", [(5, 1000, 128), (64, 5000, 896), (100, 33000, 256), (200, 2048, 192)])
def test_lm_head_argmax_epilogue_swap_ab(gemm, batch, vocab, k):
    """KX_EPI_ARGMAX, decode orientation (weights are the 128-row operand): ids must equal the argmax of the
    logits the fp32 epilogue of the same kernel writes, without [batch, vocab] ever being stored."""
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import _lib
    w, x = _rand(vocab, k, 11), _rand(batch, k, 12)
    w[vocab // 3] = w[vocab // 2]                                  # an exact tie: the lower index must win
    logits = torch.zeros(batch, vocab, device="cuda")
    gemm(w, x, 3, 1, logits)
    n_parts = 4 * ((vocab + 127) // 128)
    parts = torch.zeros(n_parts, batch, device="cuda", dtype=torch.int64)      # int2 (value bits, index) per entry
    gemm(w, x, 5, 1, parts)
    ids = torch.zeros(batch, device="cuda", dtype=torch.int32)
    L = _lib.lib()
    _lib.check(L.kx_argmax_parts(parts.data_ptr(), n_parts, batch, batch, ids.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert (ids.cpu().numpy() == _argmax_rows(logits)).all()
    _check_against_torch(ids, x, w)


@pytest.mark.parametrize("batch,vocab,k", [(256, 4096, 128), (300, 5000, 192), (512, 33000, 256)])
def test_lm_head_argmax_epilogue_token_major(gemm, batch, vocab, k):
    """Same through the CTA-pair kernel (wide decode batch): one part per feature tile."""
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import _lib
    x, w = _rand(batch, k, 13), _rand(vocab, k, 14)
    w[vocab // 3] = w[vocab // 2]
    logits = torch.zeros(batch, vocab, device="cuda")
    gemm(x, w, 3, 0, logits)
    bn = 256 if (vocab % 256 == 0 or vocab >= 2048) else 128
    n_parts = (vocab + bn - 1) // bn
    parts = torch.zeros(n_parts, batch, device="cuda", dtype=torch.int64)      # int2 (value bits, index) per entry
    gemm(x, w, 5, 0, parts)
    ids = torch.zeros(batch, device="cuda", dtype=torch.int32)
    L = _lib.lib()
    _lib.check(L.kx_argmax_parts(parts.data_ptr(), n_parts, batch, batch, ids.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert (ids.cpu().numpy() == _argmax_rows(logits)).all()
    _check_against_torch(ids, x, w)
