"""Architectures the local generator can run; random-init weights of
these shapes stand in for checkpoints (no network).  ``tiny-*`` entries are small
versions used by the parity tests."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple


@dataclass(frozen=True)
class ModelConfig:
    name: str
    family: str                 # "qwen2" | "llama"
    hidden: int
    layers: int
    n_q: int
    n_kv: int
    head_dim: int
    inter: int
    vocab: int
    tied: bool
    qkv_bias: bool
    rope_theta: float
    eps: float
    rope_scaling: Optional[Tuple[float, float, float, int]] = None   # llama3: (factor, low, high, original_max_pos)
    max_pos: int = 2048
    # std of the embedding table under the parity init (weights.py): large enough that the token's own embedding
    # stays visible in the final hidden state (top-1 logit ~10 sigma above the bulk), small enough that the layers
    # still carry most of the residual's energy.  None: the architecture is only used with the iid init.
    parity_embed_std: Optional[float] = None

    @property
    def q_dim(self) -> int:
        return self.n_q * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.n_kv * self.head_dim

    @property
    def qkv_dim(self) -> int:
        return self.q_dim + 2 * self.kv_dim

    def matmul_params(self) -> int:
        per_layer = self.hidden * self.qkv_dim + self.q_dim * self.hidden + 3 * self.hidden * self.inter
        return self.layers * per_layer

    def decode_weight_bytes(self) -> int:
        """bf16 bytes streamed per decode step (embedding rows are gathered, not streamed;
        a tied lm_head streams the embedding table once)."""
        n = self.matmul_params() + self.vocab * self.hidden
        n += self.layers * (2 * self.hidden + (self.qkv_dim if self.qkv_bias else 0)) + self.hidden
        return 2 * n

    def kv_bytes_per_token(self) -> int:
        return 2 * self.layers * self.kv_dim * 2


CONFIGS = {
    "qwen2.5-0.5b": ModelConfig("qwen2.5-0.5b", "qwen2", 896, 24, 14, 2, 64, 4864, 151_936, True, True, 1e6, 1e-6,
                                parity_embed_std=0.5),
    "llama-3.2-1b": ModelConfig("llama-3.2-1b", "llama", 2048, 16, 32, 8, 64, 8192, 128_256, True, False, 5e5, 1e-5,
                                (32.0, 1.0, 4.0, 8192), parity_embed_std=1.0),
    "llama-3.1-8b": ModelConfig("llama-3.1-8b", "llama", 4096, 32, 32, 8, 128, 14_336, 128_256, False, False, 5e5, 1e-5,
                                (8.0, 1.0, 4.0, 8192), parity_embed_std=3.0),
    # small shapes for parity tests (same code paths: GQA, bias / llama3 rope, tied / untied)
    "tiny-qwen2": ModelConfig("tiny-qwen2", "qwen2", 128, 2, 4, 2, 64, 256, 1024, True, True, 1e6, 1e-6),
    "tiny-llama": ModelConfig("tiny-llama", "llama", 256, 2, 4, 2, 64, 512, 2048, False, False, 5e5, 1e-5,
                              (8.0, 1.0, 4.0, 64), max_pos=512),
    "tiny-llama-hd128": ModelConfig("tiny-llama-hd128", "llama", 256, 2, 2, 1, 128, 512, 2048, False, False, 5e5, 1e-5,
                                    (8.0, 1.0, 4.0, 8192)),
}


def get_config(name: str) -> ModelConfig:
    return CONFIGS[name]
