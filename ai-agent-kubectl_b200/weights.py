"""Seeded random-init weights of a named architecture (there are no checkpoints offline).

Every tensor has its own generator seeded from (seed, tensor name), so a single tensor can
be regenerated on demand -- the layer-streamed CPU oracle of the 8B parity test relies on it.

Linear layers and biases: normal, std 0.02 (BASELINE / SURVEY 8d), rounded to bf16; the CPU
reference upcasts the same values.  The embedding table's std is a parameter:

  * ``embed_std=0.02`` (default): everything iid.  The residual stream is then dominated by the
    random MLP outputs and the logits are Gaussian over the vocabulary: the gap between the
    two largest of ~10^5 such logits is below the bf16 noise of the GPU path for about one
    token in five, so "32 greedy ids equal to an fp32 model" cannot hold for reasons that
    have nothing to do with the kernels.
  * ``embed_std=cfg.parity_embed_std``: the embedding is scaled up so that the token's own
    embedding keeps a cosine of roughly 10/sqrt(hidden) with the final hidden state.  With a
    tied head that alone makes the current token's logit stand ~10 sigma above the Gaussian
    bulk (top-2 margin >> bf16 noise) while the layers still contribute most of the residual's
    energy, so the logits check stays sensitive to every kernel.  Architectures with an untied
    head get the same property from ``lm_head[v + 1] = embed_direction[v]`` (the model then
    counts upwards).  This is the init the full-architecture parity tests and bench.py use
    ("choose init scales so the top-1 margin is >> bf16 noise").
"""
from __future__ import annotations

import zlib
from typing import Dict, Iterable, Optional

import torch

from .model_config import ModelConfig

LINEAR_STD = 0.02


def _gen(seed: int, name: str) -> torch.Generator:
    return torch.Generator().manual_seed((int(seed) << 32) ^ zlib.crc32(name.encode()))


def tensor_names(cfg: ModelConfig, layer: Optional[int] = None) -> Iterable[str]:
    """Names of the logical tensors: of one layer, or (layer=None) of the whole model."""
    def layer_names(i):
        out = [f"ln1.{i}", f"wq.{i}", f"wk.{i}", f"wv.{i}"]
        if cfg.qkv_bias:
            out += [f"bq.{i}", f"bk.{i}", f"bv.{i}"]
        return out + [f"wo.{i}", f"ln2.{i}", f"wg.{i}", f"wu.{i}", f"wd.{i}"]
    if layer is not None:
        return layer_names(layer)
    names = ["embed"]
    for i in range(cfg.layers):
        names += layer_names(i)
    names.append("norm")
    if not cfg.tied:
        names.append("lm_head")
    return names


def make_tensor(cfg: ModelConfig, seed: int, name: str, embed_std: float = LINEAR_STD) -> torch.Tensor:
    """One logical tensor (bf16, CPU), a pure function of (cfg, seed, name, embed_std)."""
    kind = name.split(".")[0]
    g = _gen(seed, name)

    def mat(r, c, std=LINEAR_STD):
        return (torch.randn(r, c, generator=g, dtype=torch.float32) * std).to(torch.bfloat16)

    def vec(n, base=0.0, std=LINEAR_STD):
        return (base + torch.randn(n, generator=g, dtype=torch.float32) * std).to(torch.bfloat16)

    if kind == "embed":
        return mat(cfg.vocab, cfg.hidden, embed_std)
    if kind == "lm_head":
        if embed_std == LINEAR_STD:
            return mat(cfg.vocab, cfg.hidden)
        # untied head under the parity init: row v+1 points along embedding row v (at the linear layers' scale),
        # plus an independent component, so the greedy continuation of token v is v+1 with a wide margin
        e = make_tensor(cfg, seed, "embed", embed_std).to(torch.float32) * (LINEAR_STD / embed_std)
        return (torch.roll(e, 1, dims=0) + torch.randn(cfg.vocab, cfg.hidden, generator=g) * (0.25 * LINEAR_STD)).to(torch.bfloat16)
    if kind in ("ln1", "ln2", "norm"):
        return vec(cfg.hidden, 1.0, 0.05)
    shapes = {"wq": (cfg.q_dim, cfg.hidden), "wk": (cfg.kv_dim, cfg.hidden), "wv": (cfg.kv_dim, cfg.hidden),
              "wo": (cfg.hidden, cfg.q_dim), "wg": (cfg.inter, cfg.hidden), "wu": (cfg.inter, cfg.hidden),
              "wd": (cfg.hidden, cfg.inter)}
    if kind in shapes:
        return mat(*shapes[kind])
    if kind in ("bq", "bk", "bv"):
        return vec(cfg.q_dim if kind == "bq" else cfg.kv_dim)
    raise KeyError(name)


class LazyWeights:
    """``w[name]`` regenerates the tensor from its seed every time and keeps nothing: packing an 8B-parameter model
    for the GPU then needs one tensor of host memory at a time instead of 16 GB."""

    def __init__(self, cfg: ModelConfig, seed: int, embed_std: float):
        self.cfg, self.seed, self.embed_std = cfg, seed, embed_std

    def __getitem__(self, name: str) -> torch.Tensor:
        return make_tensor(self.cfg, self.seed, name, self.embed_std)


def make_weights(cfg: ModelConfig, seed: int, embed_std: Optional[float] = None) -> Dict[str, torch.Tensor]:
    """Logical (HF-style) tensors in bf16 on the CPU:
    embed [V,H]; per layer i: ln1.i, wq.i [q_dim,H], wk.i, wv.i [kv_dim,H], (bq.i, bk.i, bv.i),
    wo.i [H,q_dim], ln2.i, wg.i, wu.i [I,H], wd.i [H,I]; norm [H]; lm_head [V,H] if untied.
    ``embed_std=None``: everything iid std 0.02 from one sequential generator (the small-model tests were
    written against these values); otherwise the per-tensor generators of make_tensor."""
    if embed_std is not None:
        return {name: make_tensor(cfg, seed, name, embed_std) for name in tensor_names(cfg)}
    g = torch.Generator().manual_seed(seed)

    def mat(r, c):
        return (torch.randn(r, c, generator=g, dtype=torch.float32) * LINEAR_STD).to(torch.bfloat16)

    def vec(n, base=0.0, std=LINEAR_STD):
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
