"""Pins the fp32 CPU decoder (oracle/cpu_model.py) to `transformers`' Qwen2 / Llama
implementations on small random-init models (architecture parity; the reference
repository itself has no model code)."""
import pytest
import torch

import kubectl_b200  # noqa: F401
from kubectl_b200 import model_config, weights
from oracle.cpu_model import RefModel


def _hf_model(cfg, w):
    tr = pytest.importorskip("transformers")
    common = dict(hidden_size=cfg.hidden, num_hidden_layers=cfg.layers, num_attention_heads=cfg.n_q,
                  num_key_value_heads=cfg.n_kv, intermediate_size=cfg.inter, vocab_size=cfg.vocab,
                  rms_norm_eps=cfg.eps, tie_word_embeddings=cfg.tied, max_position_embeddings=cfg.max_pos,
                  head_dim=cfg.head_dim, attention_dropout=0.0)
    rope = {"rope_type": "default", "rope_theta": cfg.rope_theta}
    if cfg.rope_scaling:
        f, lo, hi, orig = cfg.rope_scaling
        rope = {"rope_type": "llama3", "rope_theta": cfg.rope_theta, "factor": f, "low_freq_factor": lo,
                "high_freq_factor": hi, "original_max_position_embeddings": orig}
    if cfg.family == "qwen2":
        hc = tr.Qwen2Config(**common, rope_parameters=rope, use_sliding_window=False)
        m = tr.Qwen2ForCausalLM(hc)
    else:
        hc = tr.LlamaConfig(**common, rope_parameters=rope, attention_bias=False, mlp_bias=False)
        m = tr.LlamaForCausalLM(hc)
    sd = {"model.embed_tokens.weight": w["embed"], "model.norm.weight": w["norm"]}
    for i in range(cfg.layers):
        p = f"model.layers.{i}."
        sd[p + "input_layernorm.weight"] = w[f"ln1.{i}"]
        sd[p + "post_attention_layernorm.weight"] = w[f"ln2.{i}"]
        for hf, mine in (("q", "wq"), ("k", "wk"), ("v", "wv"), ("o", "wo")):
            sd[p + f"self_attn.{hf}_proj.weight"] = w[f"{mine}.{i}"]
        if cfg.qkv_bias:
            for hf, mine in (("q", "bq"), ("k", "bk"), ("v", "bv")):
                sd[p + f"self_attn.{hf}_proj.bias"] = w[f"{mine}.{i}"]
        for hf, mine in (("gate", "wg"), ("up", "wu"), ("down", "wd")):
            sd[p + f"mlp.{hf}_proj.weight"] = w[f"{mine}.{i}"]
    sd["lm_head.weight"] = w["embed"] if cfg.tied else w["lm_head"]
    sd = {k: v.to(torch.float32) for k, v in sd.items()}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not [k for k in missing if "rotary" not in k] and not unexpected, (missing, unexpected)
    return m.eval().to(torch.float32)


@pytest.mark.parametrize("name", ["tiny-qwen2", "tiny-llama", "tiny-llama-hd128"])
def test_matches_transformers(name):
    cfg = model_config.get_config(name)
    w = weights.make_weights(cfg, seed=3)
    ref = RefModel(cfg, w, weights.rope_table(cfg, 256))
    hf = _hf_model(cfg, w)
    g = torch.Generator().manual_seed(0)
    toks = torch.randint(0, cfg.vocab, (3, 70), generator=g)
    with torch.no_grad():
        want = hf(toks).logits
        got = ref.forward(toks, last_only=False)
    assert torch.allclose(got, want, rtol=1e-4, atol=2e-5), (got - want).abs().max()
    # incremental decoding with the KV cache reproduces the full forward
    ids, logs = ref.generate(toks[:, :40], 6, return_logits=True)
    with torch.no_grad():
        full = hf(torch.cat([toks[:, :40], ids[:, :5]], dim=1)).logits
    for t in range(6):
        assert torch.allclose(logs[t], full[:, 39 + t], rtol=1e-4, atol=2e-5)
        assert torch.equal(ids[:, t], full[:, 39 + t].argmax(-1))


def test_bf16_emulation_stays_within_the_logit_bound():
    cfg = model_config.get_config("tiny-qwen2")
    w = weights.make_weights(cfg, seed=3)
    rope = weights.rope_table(cfg, 256)
    toks = torch.randint(0, cfg.vocab, (4, 64), generator=torch.Generator().manual_seed(1))
    a = RefModel(cfg, w, rope).forward(toks)
    b = RefModel(cfg, w, rope, emulate_bf16=True).forward(toks)
    assert (a - b).abs().max() <= 1e-2 * a.abs().max()
