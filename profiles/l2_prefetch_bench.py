"""Does asking the L2 for a weight matrix ahead of time (cp.async.bulk.prefetch.L2) make the decode GEMM that reads
it faster?  For each Llama-3.2-1B decode projection at batch 64: L2 flushed / matrix prefetched and resident /
prefetch issued right before the GEMM (in flight while the GEMM starts)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import _lib  # noqa: E402

L = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64


def bench(name, feat, k, epi, splits):
    w = (torch.randn(feat, k, device="cuda") * 0.02).to(torch.bfloat16)
    x = (torch.randn(B, k, device="cuda")).to(torch.bfloat16)
    ldo = feat // 2 if epi == 2 else feat
    out = torch.zeros(B, ldo, device="cuda", dtype=torch.float32 if epi in (1, 3) else torch.bfloat16)

    def gemm():
        _lib.check(L.kx_gemm_bf16(w.data_ptr(), feat, k, x.data_ptr(), B, k, k, epi, 1, out.data_ptr(), ldo, None, splits, st))

    def timed(mode):
        ts = []
        for _ in range(7):
            flush.fill_(1)
            if mode == "hot":
                _lib.check(L.kx_l2_prefetch(w.data_ptr(), w.numel() * 2, st))
                torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            if mode == "inflight":
                _lib.check(L.kx_l2_prefetch(w.data_ptr(), w.numel() * 2, st))
            e0.record()
            gemm()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        return sorted(ts)[len(ts) // 2]
    gemm(); torch.cuda.synchronize()
    mb = w.numel() * 2 / 1e6
    c, h, f = timed("cold"), timed("hot"), timed("inflight")
    print(f"{name:<8} {mb:6.1f} MB splits {splits:2d}: cold {c:6.1f} us ({mb / c * 1e-3:5.2f} TB/s)   L2-resident {h:6.1f} us ({mb / h * 1e-3:5.2f} TB/s)"
          f"   prefetch in flight {f:6.1f} us")


for splits_scale in (1.0, 0.5):
    print(f"-- batch {B}, split factor x{splits_scale}")
    for name, feat, k, epi, s in (("qkv", 3072, 2048, 0, 8), ("wo", 2048, 2048, 1, 12), ("gate_up", 16384, 2048, 2, 2), ("down", 2048, 8192, 1, 12),
                                  ("lm_head", 128256, 2048, 3, 1)):
        bench(name, feat, k, epi, max(1, int(s * splits_scale)))
