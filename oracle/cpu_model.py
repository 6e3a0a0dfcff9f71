"""TEST INFRASTRUCTURE ONLY -- plain-PyTorch fp32 CPU decoder (Llama / Qwen2 family)
used as the oracle of the generator kernels and as the timed CPU baseline of the
generation workloads.  The product path never imports it.

Parity status: the reference repository contains no model code (its ``llm`` slot is
the OpenAI service, app.py:108-117), so for generation **parity is unpinned by the
reference**.  The pins are BASELINE.json's: first 32 greedy token ids equal, logits
within 1e-2 relative of this model with the same (bf16-rounded) weights.  The
architecture itself is pinned against ``transformers`` (Qwen2ForCausalLM /
LlamaForCausalLM, version 5.5 in this image) in tests/test_cpu_model.py.

Two numeric modes:
  * ``emulate_bf16=False`` -- everything in fp32 (the reference the 1e-2 logits
    bound is stated against).
  * ``emulate_bf16=True``  -- same fp32 arithmetic, but activations are rounded to
    bf16 exactly where the CUDA path stores bf16 (GEMM inputs, q/k/v, KV cache,
    attention output, SwiGLU output).  The small-model tests compare greedy ids
    against this mode: with iid random weights top-1 margins are tiny, and only a
    reference with the same quantisation points makes "ids equal" a statement
    about the kernels rather than about rounding luck.  DESIGN.md lists the
    rounding points.  ``deferred_decode`` (default on): single-token steps of
    at most 256 sequences round the GEMM input BEFORE the division by the rms,
    bf16(x * g), and divide the GEMM's output, as the decode kernels do
    (kx_gemm.cuh, deferred RMSNorm); tensor-parallel engines keep the plain order.
  The full-architecture tests (tests/test_parity_full_gpu.py) use the fp32 mode.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch


def _r(x: torch.Tensor, on: bool) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32) if on else x


class StreamedWeights:
    """Weights that are regenerated from their seed each time they are read and never kept:
    ``w[name]`` -> fp32 tensor (kubectl_b200.weights.make_tensor).  A forward pass reads every
    layer tensor exactly once, so an 8B-parameter fp32 model needs one layer (< 1 GB) of memory
    instead of 32 GB.  Meant for ONE teacher-forced forward over prompt + generated tokens."""

    def __init__(self, cfg, seed: int, embed_std: float):
        self.cfg, self.seed, self.embed_std = cfg, seed, embed_std

    def __getitem__(self, name: str) -> torch.Tensor:
        from kubectl_b200 import weights as wm
        return wm.make_tensor(self.cfg, self.seed, name, self.embed_std).to(torch.float32)


class RefModel:
    def __init__(self, cfg, weights, rope: torch.Tensor, emulate_bf16: bool = False, deferred_decode: bool = True):
        self.cfg, self.emu = cfg, emulate_bf16
        self.deferred_decode = deferred_decode
        self.w = weights if isinstance(weights, StreamedWeights) else {k: v.to(torch.float32) for k, v in weights.items()}
        self.rope = rope  # [max_pos, hd/2, 2]

    # -- pieces ---------------------------------------------------------------
    def _norm(self, x, g):
        var = x.pow(2).mean(-1, keepdim=True)
        return x * torch.rsqrt(var + self.cfg.eps) * g

    def _rope(self, x, pos):
        # x [B,S,heads,hd], pos [B,S]; rotate-half pairs (i, i + hd/2)
        half = self.cfg.head_dim // 2
        cs = self.rope[pos]                      # [B,S,half,2]
        cos, sin = cs[..., 0][:, :, None, :], cs[..., 1][:, :, None, :]
        x1, x2 = x[..., :half], x[..., half:]
        return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1)

    def _pre(self, x, g):
        """Input of the projections behind a norm and the factor their outputs are multiplied by afterwards."""
        B, S, _ = x.shape
        if self.emu and self.deferred_decode and S == 1 and B <= 256:
            return _r(x * g, True), torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + self.cfg.eps)
        return _r(self._norm(x, g), self.emu), 1.0

    def _layer(self, i, x, pos, cache, mask):
        c, w, emu = self.cfg, self.w, self.emu
        B, S, _ = x.shape
        h, f = self._pre(x, w[f"ln1.{i}"])
        q, k, v = (h @ w[f"wq.{i}"].T) * f, (h @ w[f"wk.{i}"].T) * f, (h @ w[f"wv.{i}"].T) * f
        if c.qkv_bias:
            q, k, v = q + w[f"bq.{i}"], k + w[f"bk.{i}"], v + w[f"bv.{i}"]
        q, k, v = _r(q, emu), _r(k, emu), _r(v, emu)
        q = _r(self._rope(q.view(B, S, c.n_q, c.head_dim), pos), emu)
        k = _r(self._rope(k.view(B, S, c.n_kv, c.head_dim), pos), emu)
        v = v.view(B, S, c.n_kv, c.head_dim)
        if cache is not None:
            pk, pv = cache[i]
            k = torch.cat([pk, k], dim=1) if pk is not None else k
            v = torch.cat([pv, v], dim=1) if pv is not None else v
            cache[i] = (k, v)
        grp = c.n_q // c.n_kv
        kk = k.repeat_interleave(grp, dim=2).permute(0, 2, 1, 3)      # [B,nq,T,hd]
        vv = v.repeat_interleave(grp, dim=2).permute(0, 2, 1, 3)
        s = (q.permute(0, 2, 1, 3) @ kk.transpose(-1, -2)) * (c.head_dim ** -0.5)
        if mask is not None:
            s = s + mask
        p = torch.softmax(s, dim=-1)
        a = _r((p @ vv).permute(0, 2, 1, 3).reshape(B, S, c.q_dim), emu)
        x = x + a @ w[f"wo.{i}"].T
        h, f = self._pre(x, w[f"ln2.{i}"])
        g, u = (h @ w[f"wg.{i}"].T) * f, (h @ w[f"wu.{i}"].T) * f
        m = _r(torch.nn.functional.silu(g) * u, emu)
        return x + m @ w[f"wd.{i}"].T

    def forward(self, tokens: torch.Tensor, pos0: int = 0, cache: Optional[list] = None,
                last_only: bool = True, last_n: Optional[int] = None) -> torch.Tensor:
        """tokens [B,S] int64 at positions pos0..pos0+S-1 -> logits [B,V] (last), [B,S,V], or
        (last_only=False, last_n=n) only the last n positions [B,n,V]."""
        B, S = tokens.shape
        x = self.w["embed"][tokens]
        pos = torch.arange(pos0, pos0 + S)[None, :].expand(B, S)
        T = pos0 + S
        mask = None
        if S > 1:
            mask = torch.full((S, T), float("-inf")).triu(diagonal=pos0 + 1)[None, None]
        for i in range(self.cfg.layers):
            x = self._layer(i, x, pos, cache, mask)
        if last_only:
            x = x[:, -1:, :]
        elif last_n is not None:
            x = x[:, -last_n:, :]
        h, f = self._pre(x, self.w["norm"])
        head = self.w["embed"] if self.cfg.tied else self.w["lm_head"]
        logits = (h @ head.T) * f
        return logits[:, 0] if last_only else logits

    @torch.no_grad()
    def generate(self, tokens: torch.Tensor, n_new: int, return_logits: bool = False
                 ) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        """Greedy decoding (temperature 0, app.py:109); ties -> lowest index like torch.argmax."""
        cache = [(None, None) for _ in range(self.cfg.layers)]
        S = tokens.shape[1]
        logits = self.forward(tokens, 0, cache)
        out, keep = [], []
        for t in range(n_new):
            nxt = logits.argmax(-1)
            out.append(nxt)
            if return_logits:
                keep.append(logits.clone())
            if t + 1 < n_new:
                logits = self.forward(nxt[:, None], S + t, cache)
        return torch.stack(out, dim=1), keep
