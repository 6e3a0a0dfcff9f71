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


def test_llama_3_2_1b_wide_decode_batch_agrees_with_narrow_batches():
    """Batches of 129-256 decode with 128-column batch tiles in the gate/up projection (one wave instead of two, no
    split-K exchange) and 64-column tiles elsewhere; the fp32 CPU model is too slow for that many sequences, so the
    check is against the same engine run on sub-batches of at most 64 (64-column tiles everywhere, pinned to the fp32
    model by the test above): same ids, logits within 1e-2 of the logit scale -- the north star's tolerance; the two tile
    shapes sum in different orders, which flips bf16 roundings of the activations, so they differ from each other about
    as much as either differs from the fp32 model (measured 3.8e-3)."""
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import engine, model_config
    cfg = model_config.get_config("llama-3.2-1b")
    B, S, n_new = 160, 32, 6
    gen = engine.Generator("llama-3.2-1b", seed=4, max_batch=B, max_seq=S + n_new, embed_std=cfg.parity_embed_std)
    toks = np.random.default_rng(5).integers(0, cfg.vocab, (B, S)).astype(np.int32)
    ids_w, log_w = gen.generate_ids(toks, n_new, want_logits=True)
    ids_g, _ = gen.generate_ids(toks, n_new)
    assert np.array_equal(ids_w, ids_g), "graph-replayed decode differs from the eager decode"
    worst = 0.0
    for b0 in range(0, B, 64):
        ids_n, log_n = gen.generate_ids(toks[b0:b0 + 64], n_new, want_logits=True)
        assert np.array_equal(ids_n, ids_w[b0:b0 + 64])
        part = log_w[:, b0:b0 + 64]
        worst = max(worst, float(((part - log_n).abs().amax(dim=(1, 2)) / log_n.abs().amax(dim=(1, 2))).max()))
    gen.close()
    print(f"wide vs narrow batch tiles: largest logit difference {worst:.2e} of the logit scale")
    assert worst <= 1e-2
