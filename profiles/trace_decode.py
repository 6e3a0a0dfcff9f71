"""In-graph timeline of one decode step (KX_DEBUG=trace: every instrumented kernel records globaltimer values).
usage: python profiles/trace_decode.py ARCH BATCH PROMPT_LEN NEW_TOKENS [layers to print]"""
import os
import sys

os.environ["KX_DEBUG"] = ",".join(filter(None, [os.environ.get("KX_DEBUG", ""), "trace"]))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import engine, model_config  # noqa: E402

arch, B, S, NEW = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
show = int(sys.argv[5]) if len(sys.argv) > 5 else 2
cfg = model_config.get_config(arch)
gen = engine.Generator(arch, seed=4, max_batch=B, max_seq=S + NEW, embed_std=cfg.parity_embed_std)
toks = np.random.default_rng(0).integers(0, 256, (B, S)).astype(np.int32)
for _ in range(3):
    gen.generate_ids(toks, NEW)
st = gen.last_stats
tr = gen.trace()
t0 = min(r[1] for r in tr)
print(f"{arch} B={B} S={S}: decode {st.decode_ms / st.decode_steps * 1e3:.1f} us/step, {len(tr)} traced launches in the last step")
print(f"{'#':>4} {'kernel':<12} {'start':>8} {'deps':>8} {'main':>8} {'depsmax':>8} {'ops/mma':>11} {'acc/xchg':>11} {'startmax':>8} {'end':>8} | {'dur':>6} {'after-deps':>10} {'gap':>6}   (us)")
prev_end = t0
per = {}
for i, (name, a, b, c, d, x4, x5, x6, x7) in enumerate(tr):
    dur, work = (d - a) / 1e3, (d - b) / 1e3
    k = per.setdefault(name, [0, 0.0, 0.0, 0.0])
    k[0] += 1; k[1] += dur; k[2] += work; k[3] += (d - max(prev_end, a)) / 1e3 if d > prev_end else 0.0
    if i < show * 8 + 2 or i >= len(tr) - 4:
        print(f"{i:>4} {name:<12} {(a - t0) / 1e3:8.2f} {(b - t0) / 1e3:8.2f} {(c - t0) / 1e3 if c else 0:8.2f} {(x4 - t0) / 1e3 if x4 else 0:8.2f} {x5 >> 32:5d}/{x5 & 0xffffffff:<5d} {x6 >> 32:5d}/{x6 & 0xffffffff:<5d} {(x7 - t0) / 1e3 if x7 else 0:8.2f} {(d - t0) / 1e3:8.2f} | "
              f"{dur:6.2f} {work:10.2f} {(a - prev_end) / 1e3:6.2f}")
    prev_end = max(prev_end, d)
print(f"span of traced launches: {(prev_end - t0) / 1e3:.1f} us")
print("per kernel: count, mean duration (start->end), mean after-deps (deps ready->end), mean exclusive time (end - max(prev end, start))")
for name, (n, dur, work, excl) in per.items():
    print(f"  {name:<12} x{n:<4} {dur / n:7.2f} {work / n:7.2f} {excl / n:7.2f}   total exclusive {excl:8.1f} us")
gen.close()
