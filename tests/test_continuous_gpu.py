"""Continuous batching on the GPU (kx_engine_session_*, include/kx.h) through the C ABI.

A slot's sequence must not depend on when its neighbours were admitted, ended or were cancelled, so the checks are
bit-exact: the same prompts through kx_engine_generate (same prefill shape, same decode batch) and through sessions
admitted all at once, staggered, with slots ending early (eos) or being released and reused mid-sequence."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

S, NEW = 64, 24


def _gen(arch, max_batch, seed=5):
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import engine
    return engine.Generator(arch, seed=seed, max_batch=max_batch, max_seq=S + NEW)


def _prompts(cfg_vocab, n, seed=11):
    return np.random.default_rng(seed).integers(0, cfg_vocab, (n, S)).astype(np.int32)


def _run_session(gen, plan, n_slots_total, steps_per_poll=3, max_steps=200):
    """plan: list of (after_step, 'admit', slot0, tokens) / (after_step, 'release', slot0, n); returns {slot: ids} of every
    slot that ended (read when it ended) and the per-slot counts at the end."""
    gen.session_open()
    out, step, k = {}, 0, 0
    plan = sorted(plan, key=lambda p: p[0])
    live = set()
    while step < max_steps:
        while k < len(plan) and plan[k][0] <= step:
            _, what, slot0, arg = plan[k]
            if what == "admit":
                gen.session_admit(slot0, arg, NEW)
                live |= set(range(slot0, slot0 + len(arg)))
            else:
                gen.session_release(slot0, arg)
                live -= set(range(slot0, slot0 + arg))
            k += 1
        if not live and k == len(plan):
            break
        if live:
            gen.session_step(steps_per_poll, max(live) + 1)
        step += steps_per_poll
        cnt, fin = gen.session_poll()
        for s in sorted(live):
            if fin[s]:
                out[s] = gen.session_read(s, 1, int(cnt[s]))[0].copy()
                gen.session_release(s, 1)
                live.discard(s)
    assert gen.session_live() == 0
    return out


@pytest.mark.parametrize("arch,n", [("tiny-qwen2", 32), ("tiny-llama", 32), ("tiny-llama", 256), ("tiny-llama", 288)])
def test_session_equals_static_batch(arch, n):
    gen = _gen(arch, n)
    toks = _prompts(gen.cfg.vocab, n)
    want, _ = gen.generate_ids(toks, NEW)
    got = _run_session(gen, [(0, "admit", 0, toks)], n)
    assert sorted(got) == list(range(n))
    for s in range(n):
        assert np.array_equal(got[s], want[s]), f"slot {s}: session {got[s]} != static batch {want[s]}"
    # and a static batch after the session still works (it ends the session)
    again, _ = gen.generate_ids(toks, NEW)
    assert np.array_equal(again, want)
    gen.close()


@pytest.mark.parametrize("arch", ["tiny-qwen2", "tiny-llama"])
def test_late_admission_does_not_disturb_running_slots(arch):
    gen = _gen(arch, 32)
    toks = _prompts(gen.cfg.vocab, 32)
    a, b = toks[:16], toks[16:]
    together = _run_session(gen, [(0, "admit", 0, a), (0, "admit", 16, b)], 32)
    staggered = _run_session(gen, [(0, "admit", 0, a), (7, "admit", 16, b)], 32)
    for s in range(32):
        assert len(staggered[s]) == NEW
        assert np.array_equal(staggered[s], together[s]), f"slot {s} depends on when its neighbours were admitted"
    gen.close()


def test_eos_ends_one_slot_and_release_frees_it_for_a_new_sequence():
    gen = _gen("tiny-llama", 32)
    toks = _prompts(gen.cfg.vocab, 32)
    ref = _run_session(gen, [(0, "admit", 0, toks)], 32)
    eos = int(ref[3][5])
    gen.set_eos(eos)
    got = _run_session(gen, [(0, "admit", 0, toks)], 32)
    ended_early = 0
    for s in range(32):
        hits = np.flatnonzero(ref[s] == eos)
        n = int(hits[0]) + 1 if hits.size else NEW
        ended_early += n < NEW
        assert len(got[s]) == n and np.array_equal(got[s], ref[s][:n]), f"slot {s}: {got[s]} vs {ref[s][:n]}"
    assert ended_early >= 1
    gen.set_eos(None)
    # cancel slots 8..15 after a few steps and reuse them for other prompts while the rest keep decoding
    new = _prompts(gen.cfg.vocab, 8, seed=99)
    alone = _run_session(gen, [(0, "admit", 8, new)], 32)
    mixed = _run_session(gen, [(0, "admit", 0, toks), (6, "release", 8, 8), (6, "admit", 8, new)], 32)
    for s in range(32):
        want = alone[s] if 8 <= s < 16 else ref[s]
        assert np.array_equal(mixed[s], want), f"slot {s} after a neighbour was cancelled and its slot reused"
    gen.close()


def test_chat_model_over_the_continuous_engine():
    import threading
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import engine
    from kubectl_b200.chat_model import B200ChatModel, ContinuousFrontend
    prompts = [f"list pods in namespace team-{i}" for i in range(32)]
    static = engine.B200Engine("tiny-qwen2", max_batch=32, prompt_len=S, max_new=16, eos=False)
    want = static.generate(prompts)
    static.close()
    m = B200ChatModel(engine=engine.B200Engine("tiny-qwen2", max_batch=32, prompt_len=S, max_new=16, eos=False, continuous=True),
                      max_wait_ms=500.0)
    assert isinstance(m.frontend, ContinuousFrontend)
    got = [o.content for o in m.batch(prompts)]
    assert got == want
    # requests that arrive while others decode are served too, each with its own completion
    res = {}
    ts = [threading.Thread(target=lambda i=i: res.__setitem__(i, m.invoke(prompts[i]).content)) for i in range(12)]
    for i, t in enumerate(ts):
        t.start()
        if i % 4 == 3:
            import time
            time.sleep(0.003)
    for t in ts:
        t.join(60)
    assert len(res) == 12 and all(isinstance(v, str) for v in res.values())
    assert m.frontend.admissions >= 2 and len(m.frontend.ttft) >= 44
    with pytest.raises(RuntimeError, match="prompt window"):
        m.invoke("x" * 500)
    m.close()


def test_service_engine_continuous_with_the_learned_template_prefix():
    """What default_engine_factory starts: the template's instruction prefix is learned from the first requests and its KV
    shared; later requests that continue it join the running sequences; one that does not waits for an idle engine,
    after which the prefix is learned again.  Completions equal the static engine's for the same prompts."""
    import threading
    import time
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import engine
    from kubectl_b200.chat_model import B200ChatModel
    head = "You are a Kubernetes CLI specialist. " * 6 + "User Request: "
    qs = [f"list pods in namespace team-{i}" for i in range(8)]
    prompts = [head + q + "\nKubectl Command:\n" for q in qs]
    kw = dict(max_batch=8, prompt_len=64, max_new=12, auto_prefix=True, max_prefix=256, eos=False)
    static = engine.B200Engine("tiny-qwen2", **kw)
    want = static.generate(prompts)
    static.close()
    eng = engine.B200Engine("tiny-qwen2", continuous=True, **kw)
    m = B200ChatModel(engine=eng, max_wait_ms=500.0)
    assert [o.content for o in m.batch(prompts)] == want
    assert eng.prefix_sets == 1
    # the same requests arriving a few at a time: same completions (the prefix is reused, slots are reused)
    res = {}
    ts = [threading.Thread(target=lambda i=i: res.__setitem__(i, m.invoke(prompts[i]).content)) for i in range(8)]
    for i, t in enumerate(ts):
        t.start()
        if i % 3 == 2:
            time.sleep(0.002)
    for t in ts:
        t.join(60)
    assert [res[i] for i in range(8)] == want and eng.prefix_sets == 1
    # a prompt with another beginning cannot join sequences that attend to the cached prefix: it is served once they are done
    other = "Translate to kubectl: " + "x" * 20
    got = m.batch([prompts[0], other, prompts[1]])
    assert got[0].content == want[0] and got[2].content == want[1]       # served with the cached prefix ...
    assert isinstance(got[1].content, str) and eng._prefix_bytes == b""   # ... which was dropped for `other` afterwards
    assert [o.content for o in m.batch(prompts)] == want and eng.prefix_sets == 2    # and learned again
    m.close()
