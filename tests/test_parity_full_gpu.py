"""Generator parity at the architectures that are benchmarked (BASELINE.json configs 3-5), full size,
against the pure-fp32 CPU model (no bf16 emulation), through the C ABI:

  1. the 32 greedy ids of every sequence equal the fp32 model's,
  2. every logit of every step within 1e-2 of the step's logit scale,
  3. gate: the fp32 model's smallest top-2 margin is at least 4x the largest GPU-vs-fp32 logit difference,
     so (1) cannot hold or fail by rounding luck (weights.py: the parity init).

oracle/parity.py states the three checks; the fp32 model is run once, teacher-forced on the GPU's tokens.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(arch, seed, B, S, n_new):
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import engine, model_config, weights
    from oracle import parity
    from oracle.cpu_model import RefModel
    cfg = model_config.get_config(arch)
    w = weights.make_weights(cfg, seed, embed_std=cfg.parity_embed_std)
    gen = engine.Generator(arch, weights=w, max_batch=B, max_seq=S + n_new)
    toks = torch.randint(0, cfg.vocab, (B, S), generator=torch.Generator().manual_seed(seed + 100))
    ids, logits = gen.generate_ids(toks.numpy().astype(np.int32), n_new, want_logits=True)
    ids_graph, _ = gen.generate_ids(toks.numpy().astype(np.int32), n_new)       # the CUDA-graph decode loop bench.py times
    gen.close()
    assert np.array_equal(ids, ids_graph), "graph-replayed decode differs from the eager decode"
    ref = RefModel(cfg, w, weights.rope_table(cfg, cfg.max_pos), emulate_bf16=False)
    rep = parity.teacher_forced_report(ref, toks, ids, logits.cpu())
    print(f"{arch}: {rep}")
    parity.assert_parity(rep, arch)
    assert len({tuple(r) for r in ids.tolist()}) == B                            # the sequences are not all alike


def test_llama_3_2_1b_headline_architecture():
    """BASELINE configs[3] (the bench default): Llama-3.2-1B, vocabulary 128,256, llama3 rope scaling, tied head;
    256-token prompts, 32 greedy tokens, on as many sequences as the CPU model finishes in seconds."""
    _run("llama-3.2-1b", seed=4, B=4, S=256, n_new=32)


def test_qwen25_05b_config3_architecture():
    """BASELINE configs[2]: Qwen2.5-0.5B (151,936 vocabulary, GQA 14/2, q/k/v biases), 256 + 32 tokens."""
    _run("qwen2.5-0.5b", seed=3, B=4, S=256, n_new=32)
