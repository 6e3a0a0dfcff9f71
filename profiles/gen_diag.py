"""Numerics probe: GPU generator vs the CPU oracle (bf16-emulating and pure fp32)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kubectl_b200  # noqa
from kubectl_b200 import engine, model_config, weights
from oracle.cpu_model import RefModel
arch = sys.argv[1] if len(sys.argv) > 1 else "qwen2.5-0.5b"
B, S, N, seed = 4, 256, 32, int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.set_num_threads(os.cpu_count())
cfg = model_config.get_config(arch)
w = weights.make_weights(cfg, seed)
rope = weights.rope_table(cfg, cfg.max_pos)
gen = engine.Generator(arch, weights=w, max_batch=B, max_seq=S + N)
toks = torch.randint(0, 256, (B, S), generator=torch.Generator().manual_seed(seed + 100))
ids, logits = gen.generate_ids(toks.numpy().astype(np.int32), N, want_logits=True)
logits = logits.cpu()
for name, emu in (("emu", True), ("fp32", False)):
    ref = RefModel(cfg, w, rope, emulate_bf16=emu)
    rid, rlog = ref.generate(toks, N, return_logits=True)
    rid = rid.numpy()
    alive = np.ones(B, bool)
    for t in range(N):
        mx = rlog[t].abs().max().item()
        err = (logits[t][alive] - rlog[t][alive]).abs().max().item() if alive.any() else float("nan")
        top2 = rlog[t].topk(2, dim=-1).values
        margin = (top2[:, 0] - top2[:, 1]).numpy()
        flips = [(b, float(margin[b])) for b in range(B) if alive[b] and ids[b, t] != rid[b, t]]
        print(f"{name} t={t:2d} alive={alive.sum()} max|logit|={mx:.3f} err={err:.2e} rel={err / mx:.2e} "
              f"min_margin={margin[alive].min() if alive.any() else float('nan'):.3e} flips={flips}")
        for b, _ in flips:
            alive[b] = False
