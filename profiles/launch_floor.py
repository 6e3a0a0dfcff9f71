"""Per-kernel floor of the decode GEMM inside a CUDA graph: python profiles/launch_floor.py
Captures a chain of N identical kx_gemm_bf16 (swap-AB, cluster split-K, PDL) launches into one graph and
times replays: us per launch for a shape whose weights are tiny (pure latency) and for real decode shapes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import _lib  # noqa: E402

L = _lib.lib()
N = 200
for name, feat, k, B, splits, epi in (("tiny 128x64", 128, 64, 64, 1, 0), ("tiny split 8", 128, 512, 64, 8, 0),
                                       ("qwen wo", 896, 896, 64, 14, 1), ("qwen qkv", 1152, 896, 64, 14, 0),
                                       ("llama wo", 2048, 2048, 64, 12, 1), ("llama qkv", 3072, 2048, 64, 8, 0)):
    w = (torch.randn(feat, k, device="cuda") * 0.05).to(torch.bfloat16)
    x = (torch.randn(B, k, device="cuda") * 0.5).to(torch.bfloat16)
    out = torch.zeros(B, feat, device="cuda", dtype=torch.float32 if epi == 1 else torch.bfloat16)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        st = s.cuda_stream
        def run():
            _lib.check(L.kx_gemm_bf16(w.data_ptr(), feat, k, x.data_ptr(), B, k, k, epi, 1, out.data_ptr(), feat, None, splits, st))
        run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(N):
                run()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(5):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
        print(f"{name:14s} feat={feat:5d} k={k:5d} splits={splits:2d}: {e0.elapsed_time(e1) * 1e3 / (5 * N):6.2f} us per launch in a graph", flush=True)
