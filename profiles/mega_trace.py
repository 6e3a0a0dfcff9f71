"""Phase timeline of the persistent decode kernel: KX_MEGA_TRACE=1 python profiles/mega_trace.py ARCH BATCH PROMPT
Prints, for CTA 0, the time between consecutive barrier events of the last decode step (tag = 2*phase on
arrival, 2*phase+1 when the wait for that phase returned; phase = 7*layer + k)."""
import os
import sys

os.environ.setdefault("KX_MEGA_TRACE", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import engine  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
B, S = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 64), (3, 256)))
gen = engine.Generator(arch, seed=3, max_batch=B, max_seq=S + 8)
toks = np.random.default_rng(0).integers(0, 256, (B, S)).astype(np.int32)
gen.generate_ids(toks, 6)
tr = gen.mega_trace(4096)
names = {0: "norm1", 1: "qkv", 2: "attn", 3: "wo", 4: "norm2", 5: "wgu", 6: "wd"}
t0 = tr[0][1]
prev = t0
per = {}
for tag, ns in tr[1:]:
    ph, kind = tag // 2, "wait" if tag & 1 else "arrive"
    layer, k = divmod(ph, 7)
    per.setdefault((names[k], kind), []).append((ns - prev) / 1e3)
    if layer in (0, 1, 7):
        print(f"layer {layer:2d} {names[k]:6s} {kind:7s} +{(ns - prev) / 1e3:7.2f} us  (t = {(ns - t0) / 1e3:8.2f} us)")
    prev = ns
print("mean us between events, by kind:")
for k, v in per.items():
    print(f"  {k[0]:6s} {k[1]:7s} {np.mean(v):7.2f}  (n={len(v)})")
print(f"total {(tr[-1][1] - t0) / 1e3:.1f} us for {len(tr)} events")
