"""Host logic of continuous batching (chat_model.ContinuousFrontend) against a slot engine that lives on the CPU:
slot reuse, joining at a step boundary, in-flight cancellation, single flight, held requests, per-request errors.
The GPU engine behind the same protocol is covered by tests/test_continuous_gpu.py."""
import threading
import time

import numpy as np
import pytest

from kubectl_b200.chat_model import B200ChatModel, ContinuousFrontend, SlotEngine, _runs


class FakeSlots(SlotEngine):
    """Completion of prompt p is 'kubectl get ' + p, one character per decode step."""

    def __init__(self, max_batch=4, step_sleep=0.0):
        self.max_batch, self.step_sleep = max_batch, step_sleep
        self.text = [None] * max_batch
        self.cnt = np.full(max_batch, -1, np.int32)
        self.log = []
        self.fail_next_step = False
        self.lock = threading.Lock()

    def slots_open(self):
        self.log.append(("open",))
        self.text = [None] * self.max_batch
        self.cnt[:] = -1

    def slots_prepare(self, prompts, idle):
        out = []
        for p in prompts:
            if p.startswith("TOOLONG"):
                out.append(ValueError("prompt too long"))
            elif p.startswith("other-prefix") and not idle:
                out.append(None)
            else:
                out.append(np.frombuffer(p.encode(), np.uint8))
        return out

    def slots_admit(self, slot0, rows):
        live = [s for s in range(self.max_batch) if self.cnt[s] >= 0]
        self.log.append(("admit", slot0, len(rows), tuple(live)))
        for j, r in enumerate(rows):
            assert self.cnt[slot0 + j] < 0, "admitted into a live slot"
            self.text[slot0 + j] = "kubectl get " + bytes(r).decode()
            self.cnt[slot0 + j] = 1

    def slots_step(self, n_steps, n_slots):
        if self.fail_next_step:
            self.fail_next_step = False
            raise MemoryError("device fell over")
        if self.step_sleep:
            time.sleep(self.step_sleep)
        for s in range(n_slots):
            if self.cnt[s] >= 0:
                self.cnt[s] = min(len(self.text[s]), self.cnt[s] + n_steps)

    def slots_poll(self):
        fin = np.array([1 if (c >= 0 and c >= len(t or "")) else 0 for c, t in zip(self.cnt, self.text)], np.int32)
        return self.cnt.copy(), fin

    def slots_read(self, slot0, counts):
        return [self.text[slot0 + j][:c] for j, c in enumerate(counts)]

    def slots_release(self, slot0, n):
        self.log.append(("release", slot0, n))
        for s in range(slot0, slot0 + n):
            self.cnt[s] = -1
            self.text[s] = None


def _model(eng, **kw):
    m = B200ChatModel(engine=eng, **kw)
    assert isinstance(m.frontend, ContinuousFrontend)
    return m


def test_runs():
    assert _runs([3, 0, 1, 5, 6, 7]) == [(0, 2), (3, 1), (5, 3)]
    assert _runs([]) == []


def test_more_requests_than_slots_are_all_answered():
    eng = FakeSlots(max_batch=4)
    m = _model(eng)
    prompts = [f"pods-{i}" * (1 + i % 3) for i in range(11)]
    outs = m.batch(prompts)
    assert [o.content for o in outs] == ["kubectl get " + p for p in prompts]
    assert sum(e[2] for e in eng.log if e[0] == "admit") == 11
    assert all(c < 0 for c in eng.cnt)                       # every slot was given back
    assert len(m.frontend.ttft) == 11 and len(m.frontend.latency) == 11
    m.close()


def test_a_late_request_joins_the_running_sequences_and_leaves_first():
    eng = FakeSlots(max_batch=4, step_sleep=0.002)
    m = _model(eng)
    res = {}

    def call(name, prompt):
        res[name] = (m.invoke(prompt).content, time.monotonic())

    long_t = threading.Thread(target=call, args=("long", "x" * 600))
    long_t.start()
    time.sleep(0.05)                                         # the long one is decoding by now
    short_t = threading.Thread(target=call, args=("short", "ns"))
    short_t.start()
    long_t.join(10); short_t.join(10)
    assert res["short"][0] == "kubectl get ns" and res["long"][0] == "kubectl get " + "x" * 600
    assert res["short"][1] < res["long"][1]                  # answered while the long sequence was still running
    admits = [e for e in eng.log if e[0] == "admit"]
    assert len(admits) == 2 and admits[1][3] == (0,)         # admitted next to the live slot 0
    m.close()


def test_a_request_whose_waiter_left_frees_its_slot_mid_sequence():
    eng = FakeSlots(max_batch=2, step_sleep=0.005)
    m = _model(eng, request_timeout=0.05)
    with pytest.raises(Exception):
        m.invoke("y" * 5000)                                 # cannot finish within the deadline
    t0 = time.monotonic()
    while m.frontend.cancelled_in_flight == 0 and time.monotonic() - t0 < 5:
        time.sleep(0.01)
    assert m.frontend.cancelled_in_flight == 1
    assert ("release", 0, 1) in eng.log and eng.cnt[0] < 0
    m.request_timeout = None
    assert m.invoke("ns").content == "kubectl get ns"        # the engine carries on
    m.close()


def test_identical_prompts_share_one_slot_also_with_a_running_sequence():
    eng = FakeSlots(max_batch=4, step_sleep=0.002)
    m = _model(eng)
    res = []
    ts = [threading.Thread(target=lambda: res.append(m.invoke("z" * 300).content)) for _ in range(2)]
    ts[0].start()
    time.sleep(0.05)
    ts[1].start()
    for t in ts:
        t.join(10)
    assert res == ["kubectl get " + "z" * 300] * 2
    assert sum(e[2] for e in eng.log if e[0] == "admit") == 1 and m.frontend.coalesced == 1
    outs = m.batch(["a", "a", "b"])
    assert [o.content for o in outs] == ["kubectl get a", "kubectl get a", "kubectl get b"]
    m.close()


def test_a_held_request_waits_for_an_idle_engine_and_nothing_overtakes_it():
    eng = FakeSlots(max_batch=4, step_sleep=0.002)
    m = _model(eng)
    res = {}

    def call(name, prompt):
        res[name] = m.invoke(prompt).content

    ts = [threading.Thread(target=call, args=("long", "w" * 200))]
    ts[0].start()
    time.sleep(0.05)
    ts.append(threading.Thread(target=call, args=("held", "other-prefix q")))
    ts[1].start()
    time.sleep(0.02)
    ts.append(threading.Thread(target=call, args=("after", "later")))
    ts[2].start()
    for t in ts:
        t.join(10)
    assert res == {"long": "kubectl get " + "w" * 200, "held": "kubectl get other-prefix q", "after": "kubectl get later"}
    admits = [e for e in eng.log if e[0] == "admit"]
    assert admits[1][3] == ()                                # the held request was admitted into an idle engine
    assert len(admits) == 2 or admits[2][3] != (0,)          # 'after' did not join the long sequence ahead of it
    m.close()


def test_one_bad_prompt_fails_alone_and_an_engine_failure_is_a_runtime_error_for_everyone():
    eng = FakeSlots(max_batch=4)
    m = _model(eng)
    with pytest.raises(RuntimeError, match="too long"):
        m.invoke("TOOLONG ...")
    assert m.invoke("ok").content == "kubectl get ok"
    eng.fail_next_step = True
    with pytest.raises(RuntimeError, match="MemoryError"):
        m.invoke("boom")
    assert m.invoke("again").content == "kubectl get again"  # the session was opened again
    assert [e for e in eng.log if e[0] == "open"].__len__() == 2
    m.close()


# ---- B200Engine's admission rules (engine.py: slots_prepare), with the GPU generator replaced by a recorder ----
class _GenStub:
    def __init__(self):
        self.calls = []

    def set_prefix(self, toks):
        self.calls.append(("set_prefix", len(toks)))

    def session_open(self):
        self.calls.append(("session_open",))


def _engine(**kw):
    from kubectl_b200 import engine
    e = object.__new__(engine.B200Engine)                    # no GPU: only the host-side prompt handling is under test
    e.gen = _GenStub()
    e.on_overflow = kw.get("on_overflow", "error")
    e.auto_prefix, e.max_prefix = kw.get("auto_prefix", True), 256
    e._prefix_bytes, e.prefix_sets = b"", 0
    e.prompt_len, e.max_new, e.max_batch = 64, 8, 8
    return e


HEAD = "You are a Kubernetes CLI specialist. " * 4 + "User Request: "


def test_prepare_learns_the_prefix_when_idle_and_keeps_it_while_requests_continue_it():
    from kubectl_b200 import engine, tokenizer
    e = _engine()
    prompts = [HEAD + q for q in ("list pods", "get nodes", "show services in ns1")]
    rows = e.slots_prepare(prompts, idle=True)
    assert all(isinstance(r, np.ndarray) and r.shape == (64,) for r in rows)
    assert e.prefix_sets == 1 and e._prefix_bytes and HEAD.encode().startswith(e._prefix_bytes[:len(HEAD) - 16])
    assert ("session_open",) in e.gen.calls                  # set_prefix ended the session: it is opened again
    n_calls = len(e.gen.calls)
    # busy engine: a prompt that continues the prefix joins, one that does not is held (None), a long one fails alone
    rows = e.slots_prepare([HEAD + "describe pod x", "Translate: y", HEAD + "z" * 200], idle=False)
    assert isinstance(rows[0], np.ndarray) and rows[1] is None and isinstance(rows[2], engine.PromptTooLong)
    assert len(e.gen.calls) == n_calls                       # nothing touched the prefix while sequences depend on it
    # the suffix is what follows the cached prefix, right-aligned in the window
    tail = (HEAD + "describe pod x").encode()[len(e._prefix_bytes):]
    assert bytes(rows[0][-len(tail):].astype(np.uint8)) == tail and rows[0][0] == 32
    # idle again, but somebody still continues the prefix: it stays, the stranger keeps waiting
    rows = e.slots_prepare([HEAD + "get pods", "Translate: y"], idle=True)
    assert isinstance(rows[0], np.ndarray) and rows[1] is None and e.prefix_sets == 1
    # idle and only strangers: the prefix is learned again from them (here: dropped, they share nothing page-sized)
    rows = e.slots_prepare(["Translate: y"], idle=True)
    assert isinstance(rows[0], np.ndarray) and rows[0][0] == tokenizer.BOS and e._prefix_bytes == b""


def test_prepare_without_a_prefix_cache_checks_the_window():
    from kubectl_b200 import engine, tokenizer
    e = _engine(auto_prefix=False)
    rows = e.slots_prepare(["kubectl get pods", "x" * 64], idle=False)
    assert rows[0][0] == tokenizer.BOS and bytes(rows[0][-16:].astype(np.uint8)) == b"kubectl get pods"
    assert isinstance(rows[1], engine.PromptTooLong)
    e = _engine(auto_prefix=False, on_overflow="truncate")
    rows = e.slots_prepare(["y" * 100], idle=True)
    assert rows[0][0] == tokenizer.BOS and bytes(rows[0][1:].astype(np.uint8)) == b"y" * 63
