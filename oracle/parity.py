"""TEST INFRASTRUCTURE ONLY -- the generator's parity gate (BASELINE.json north_star checks 1-2)
stated once, for tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.

The checks, against the fp32 CPU model (oracle/cpu_model.py, emulate_bf16=False) with the same weights:
  1. greedy ids equal for every generated token.  Verified by teacher forcing: the fp32 model is run ONCE
     over prompt + the GPU's tokens and must pick, at every generated position, the token the GPU picked.
     By induction that is exactly "the fp32 model's own greedy continuation equals the GPU's".
  2. logits within 1e-2 of the logit scale (largest |logit| of the step) for every logit of every step.
  3. the gate that makes (1) a statement about the kernels rather than about rounding luck: the smallest top-2
     margin of the fp32 logits must be at least 4x the largest GPU-vs-fp32 logit difference (weights.py explains
     the init that provides such margins).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

LOGIT_TOL = 1e-2          # north_star: "logits within 1e-2 relative"
MARGIN_OVER_NOISE = 4.0   # gate (3)


@torch.no_grad()
def teacher_forced_report(ref_f32, prompt: torch.Tensor, gpu_ids: np.ndarray,
                          gpu_logits: Optional[torch.Tensor] = None) -> Dict:
    """ref_f32: RefModel(emulate_bf16=False); prompt int64 [B, S] (cached prefix included); gpu_ids int [B, n];
    gpu_logits fp32 [n, B, V] on the CPU (optional).  Returns the numbers the gate is stated on."""
    B, S = prompt.shape
    n = gpu_ids.shape[1]
    ids = torch.from_numpy(np.asarray(gpu_ids).astype(np.int64))
    hist = torch.cat([prompt, ids[:, :-1]], dim=1)
    ref = ref_f32.forward(hist, last_only=False, last_n=n)            # [B, n, V]: logits that choose token t
    ref_ids = ref.argmax(-1)
    differs = (ref_ids != ids).numpy()
    first_diff = [int(np.argmax(np.append(differs[b], True))) for b in range(B)]
    top2 = ref.topk(2, dim=-1).values
    margin = top2[..., 0] - top2[..., 1]                               # [B, n]
    scale = ref.abs().amax(dim=(0, 2))                                 # [n] logit scale of each step
    out = {"first_differing_token": first_diff, "new_tokens": n,
           "min_margin_over_scale": float((margin / scale[None, :]).min())}
    if gpu_logits is not None:
        err = (gpu_logits.permute(1, 0, 2) - ref).abs().amax(dim=(0, 2))     # [n] largest difference of each step
        out["max_logit_err_over_scale"] = float((err / scale).max())
        out["min_margin_over_noise"] = float((margin.amin(dim=0) / err).min())
    return out


def assert_parity(rep: Dict, what: str = "") -> None:
    n = rep["new_tokens"]
    assert all(f == n for f in rep["first_differing_token"]), \
        f"{what}: greedy ids differ from the fp32 oracle, first differing token per sequence {rep['first_differing_token']}"
    if "max_logit_err_over_scale" in rep:
        assert rep["max_logit_err_over_scale"] <= LOGIT_TOL, \
            f"{what}: logits off by {rep['max_logit_err_over_scale']:.2e} of the logit scale (bound {LOGIT_TOL})"
        assert rep["min_margin_over_noise"] >= MARGIN_OVER_NOISE, \
            f"{what}: top-2 margin only {rep['min_margin_over_noise']:.1f}x the bf16 noise (gate {MARGIN_OVER_NOISE}x)"
