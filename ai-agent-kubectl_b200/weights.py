"""Seeded random-init weights of a named architecture (BASELINE: normal, std 0.02,
then rounded to bf16; the CPU reference upcasts the same values)."""
from __future__ import annotations

from typing import Dict

import torch

from .model_config import ModelConfig


def make_weights(cfg: ModelConfig, seed: int) -> Dict[str, torch.Tensor]:
    """Logical (HF-style) tensors in bf16 on the CPU:
    embed [V,H]; per layer i: ln1.i, wq.i [q_dim,H], wk.i, wv.i [kv_dim,H], (bq.i, bk.i, bv.i),
    wo.i [H,q_dim], ln2.i, wg.i, wu.i [I,H], wd.i [H,I]; norm [H]; lm_head [V,H] if untied."""
    g = torch.Generator().manual_seed(seed)

    def mat(r, c):
        return (torch.randn(r, c, generator=g, dtype=torch.float32) * 0.02).to(torch.bfloat16)

    def vec(n, base=0.0, std=0.02):
        return (base + torch.randn(n, generator=g, dtype=torch.float32) * std).to(torch.bfloat16)

    w = {"embed": mat(cfg.vocab, cfg.hidden)}
    for i in range(cfg.layers):
        w[f"ln1.{i}"] = vec(cfg.hidden, 1.0, 0.05)
        w[f"wq.{i}"] = mat(cfg.q_dim, cfg.hidden)
        w[f"wk.{i}"] = mat(cfg.kv_dim, cfg.hidden)
        w[f"wv.{i}"] = mat(cfg.kv_dim, cfg.hidden)
        if cfg.qkv_bias:
            w[f"bq.{i}"] = vec(cfg.q_dim)
            w[f"bk.{i}"] = vec(cfg.kv_dim)
            w[f"bv.{i}"] = vec(cfg.kv_dim)
        w[f"wo.{i}"] = mat(cfg.hidden, cfg.q_dim)
        w[f"ln2.{i}"] = vec(cfg.hidden, 1.0, 0.05)
        w[f"wg.{i}"] = mat(cfg.inter, cfg.hidden)
        w[f"wu.{i}"] = mat(cfg.inter, cfg.hidden)
        w[f"wd.{i}"] = mat(cfg.hidden, cfg.inter)
    w["norm"] = vec(cfg.hidden, 1.0, 0.05)
    if not cfg.tied:
        w["lm_head"] = mat(cfg.vocab, cfg.hidden)
    return w


def rope_inv_freq(cfg: ModelConfig) -> torch.Tensor:
    """Per-pair inverse frequencies [head_dim/2] in fp32, with the llama3 rescale
    (transformers modeling_rope_utils._compute_llama3_parameters) when configured."""
    import math
    half = cfg.head_dim // 2
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, half, dtype=torch.float64) * 2.0 / cfg.head_dim))
    if cfg.rope_scaling is not None:
        factor, low, high, orig = cfg.rope_scaling
        low_wl, high_wl = orig / low, orig / high
        wl = 2 * math.pi / inv
        smooth = (orig / wl - low) / (high - low)
        scaled = torch.where(wl > low_wl, inv / factor, inv)
        mid = (1 - smooth) * inv / factor + smooth * inv
        is_mid = (wl <= low_wl) & (wl >= high_wl)
        inv = torch.where(is_mid, mid, scaled)
    return inv.to(torch.float32)


def rope_table(cfg: ModelConfig, max_pos: int) -> torch.Tensor:
    """[max_pos, head_dim/2, 2] fp32 (cos, sin)."""
    inv = rope_inv_freq(cfg)
    ang = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv[None, :]
    return torch.stack([ang.cos(), ang.sin()], dim=-1).contiguous()
