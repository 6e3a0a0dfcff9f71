"""Decode-GEMM sweep: python profiles/sweep_decode_gemm.py
For each (features, k) weight shape of the named architectures and each batch, times
kx_gemm_bf16(swap=1) over the split-K factors, rotating over enough weight copies that no
launch finds its weights in the 126 MB L2.  Prints us per launch and weight GB/s."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import _lib  # noqa: E402

L = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
SHAPES = {  # name: (features, k, epilogue)
    "1b.qkv": (3072, 2048, 0), "1b.wo": (2048, 2048, 1), "1b.wgu": (16384, 2048, 2), "1b.wd": (2048, 8192, 1),
    "8b.qkv": (6144, 4096, 0), "8b.wo": (4096, 4096, 1), "8b.wgu": (28672, 4096, 2), "8b.wd": (4096, 14336, 1),
    "8b/2.qkv": (3072, 4096, 0), "8b/2.wo": (4096, 2048, 1), "8b/2.wgu": (14336, 4096, 2), "8b/2.wd": (4096, 7168, 1),
    "q.qkv": (1152, 896, 0), "q.wo": (896, 896, 1), "q.wgu": (9728, 896, 2), "q.wd": (896, 4864, 1),
}
SPLITS = tuple(int(x) for x in os.environ.get("SPLITS", "1,2,3,4,5,6,7,8").split(","))
only = sys.argv[1:] or list(SHAPES)
batches = [64]
for name in only:
    feat, k, epi = SHAPES[name]
    wbytes = feat * k * 2
    copies = max(2, min(16, (300 << 20) // wbytes + 1))
    w = [(torch.randn(feat, k, device="cuda") * 0.05).to(torch.bfloat16) for _ in range(copies)]
    for B in batches:
        x = (torch.randn(B, k, device="cuda") * 0.5).to(torch.bfloat16)
        ldo = feat // 2 if epi == 2 else feat
        out = torch.zeros(B, ldo, device="cuda", dtype=torch.float32 if epi in (1, 3) else torch.bfloat16)
        line = []
        for splits in SPLITS:
            def run(i):
                _lib.check(L.kx_gemm_bf16(w[i % copies].data_ptr(), feat, k, x.data_ptr(), B, k, k, epi, 1,
                                          out.data_ptr(), ldo, None, splits, st))
            for i in range(copies):
                run(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 5 * copies
            e0.record()
            for i in range(n):
                run(i)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / n
            line.append(f"s{splits}:{us:6.1f}us {wbytes / us / 1e3:5.0f}GB/s")
        tiles = (feat + 127) // 128
        print(f"{name:9s} B={B:3d} tiles={tiles:3d} kb={k // 64:3d} {wbytes / 1e6:6.1f}MB | " + " | ".join(line), flush=True)
