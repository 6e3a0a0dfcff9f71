"""Generation timing probe: python profiles/run_gen.py ARCH BATCH PROMPT NEW [reps]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import engine, model_config  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "qwen2.5-0.5b"
B, S, N = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 64), (3, 256), (4, 32)))
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
cfg = model_config.get_config(arch)
t0 = time.time()
gen = engine.Generator(arch, seed=3, max_batch=B, max_seq=S + N)
print(f"engine up in {time.time() - t0:.1f}s, device bytes {gen.device_bytes() / 1e9:.2f} GB")
toks = np.random.default_rng(0).integers(0, 256, (B, S)).astype(np.int32)
for r in range(reps):
    ids, _ = gen.generate_ids(toks, N)
    s = gen.last_stats
    flops = 2.0 * cfg.matmul_params() * B * S
    wbytes = cfg.decode_weight_bytes()
    kvb = B * (S + N / 2) * cfg.kv_bytes_per_token()
    step_ms = s.decode_ms / max(1, s.decode_steps)
    print(f"rep {r}: prefill {s.prefill_ms:.2f} ms ({flops / s.prefill_ms / 1e9:.0f} TFLOP/s linear), "
          f"decode {s.decode_ms:.2f} ms / {s.decode_steps} steps = {step_ms:.3f} ms/step "
          f"({(wbytes + kvb) / step_ms / 1e6:.0f} GB/s), launches {s.launches}, "
          f"{B / ((s.prefill_ms + s.decode_ms) / 1e3):.0f} commands/s")
