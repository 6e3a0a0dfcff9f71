"""Shared-prefix generation probe: python profiles/run_prefix.py ARCH BATCH PREFIX SUFFIX NEW [reps]
(KX_DEBUG=no-shared-prefix in the environment makes every sequence re-read the prefix K/V, for A/B timing)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import engine, model_config  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "llama-3.1-8b"
B, P, S, N = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 64), (3, 1024), (4, 32), (5, 64)))
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 3
cfg = model_config.get_config(arch)
t0 = time.time()
gen = engine.Generator(arch, seed=5, max_batch=B, max_seq=P + S + N)
rng = np.random.default_rng(0)
gen.set_prefix(rng.integers(0, 256, P).astype(np.int32))
print(f"engine + prefix up in {time.time() - t0:.1f}s, device bytes {gen.device_bytes() / 1e9:.2f} GB, "
      f"KX_DEBUG={os.environ.get('KX_DEBUG', '')}")
toks = rng.integers(0, 256, (B, S)).astype(np.int32)
for r in range(reps):
    ids, _ = gen.generate_ids(toks, N)
    s = gen.last_stats
    step_ms = s.decode_ms / max(1, s.decode_steps)
    print(f"rep {r}: prefill {s.prefill_ms:.2f} ms, decode {step_ms:.3f} ms/step, launches {s.launches}, "
          f"{B / ((s.prefill_ms + s.decode_ms) / 1e3):.1f} commands/s, ids checksum {int(ids.astype(np.int64).sum())}")
