"""RMSNorm (k5) by itself through the C ABI (kx_rmsnorm) against a plain PyTorch fp32 reference of the op: the row
kernel (decode / small M, with and without a row map and pending split-K partial sums) and the warp-per-row kernel
(prefill).  Tolerance: the output is bf16, so one bf16 ulp (2^-8 relative) of the fp32 reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _norm(x, g, y, eps, row_map=None, part=None, splits=0, stride=0):
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import _lib
    _lib.check(_lib.lib().kx_rmsnorm(x.data_ptr(), g.data_ptr(), y.data_ptr(), y.shape[0], x.shape[1], eps,
                                     None if row_map is None else row_map.data_ptr(),
                                     None if part is None else part.data_ptr(), splits, stride,
                                     torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()


def _ref(x, g, eps):
    x = x.float()
    return x * torch.rsqrt((x * x).mean(dim=1, keepdim=True) + eps) * g.float()


def _close(y, want):
    err = (y.float() - want).abs()
    assert (err <= want.abs() * 2.0 ** -8 + 1e-6).all(), float((err / (want.abs() + 1e-6)).max())


@pytest.mark.parametrize("M,H", [(1, 128), (64, 896), (512, 2048), (300, 4096), (2048, 1024), (5000, 2048), (33000, 896)])
def test_rmsnorm_matches_torch(M, H):
    g = torch.Generator(device="cuda").manual_seed(M + H)
    x = torch.randn(M, H, generator=g, device="cuda") * 3.0
    gain = (1.0 + 0.05 * torch.randn(H, generator=g, device="cuda")).to(torch.bfloat16)
    y = torch.zeros(M, H, device="cuda", dtype=torch.bfloat16)
    _norm(x, gain, y, 1e-5)
    _close(y, _ref(x, gain, 1e-5))


def test_rmsnorm_row_map_and_pending_partials():
    M, H, S = 40, 2048, 4
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(3 * M, H, generator=g, device="cuda")
    gain = (1.0 + 0.05 * torch.randn(H, generator=g, device="cuda")).to(torch.bfloat16)
    rows = torch.randperm(3 * M, generator=g, device="cuda")[:M].to(torch.int32)
    y = torch.zeros(M, H, device="cuda", dtype=torch.bfloat16)
    _norm(x, gain, y, 1e-6, row_map=rows)
    _close(y, _ref(x[rows.long()], gain, 1e-6))
    # residual += sum of the slices (in slice order), written back, then normalised
    x2 = torch.randn(M, H, generator=g, device="cuda")
    part = torch.randn(S, M, H, generator=g, device="cuda")
    want_x = x2.clone()
    acc = part[0].clone()
    for s in range(1, S):
        acc += part[s]
    want_x += acc
    _norm(x2, gain, y, 1e-6, part=part, splits=S, stride=M * H)
    assert torch.equal(x2, want_x)
    _close(y, _ref(want_x, gain, 1e-6))
