"""Replica router: host logic only, CPU."""
import threading
import time
import zlib

import pytest

import kubectl_b200  # noqa: F401
from kubectl_b200 import chat_model, router, synth


def _crc_keys(prompts):
    return [zlib.crc32(synth.query_of_prompt(p).encode()) for p in prompts]


def _prompt(q):
    return "You are a kubectl expert.\nUser Request: " + q + "\nKubectl Command:\n"


def test_affinity_sends_equal_requests_to_one_replica_and_keeps_order():
    engines = [chat_model.CallableEngine(lambda p, i=i: f"r{i}:" + synth.query_of_prompt(p), max_batch=4) for i in range(3)]
    r = router.ReplicaRouter(engines, key_fn=_crc_keys)
    qs = [f"list pods in ns{i % 7}" for i in range(40)]
    outs = r.generate([_prompt(q) for q in qs])
    assert [o.split(":", 1)[1] for o in outs] == qs                       # order preserved
    seen = {}
    for q, o in zip(qs, outs):
        assert seen.setdefault(q, o.split(":")[0]) == o.split(":")[0]     # the same query always meets the same replica
    assert sum(r.routed) == 40 and r.in_flight == [0, 0, 0]
    assert all(b <= 4 for e in engines for b in e.batches)                # max_batch of a replica respected
    r.close()


def test_least_loaded_balances_and_runs_replicas_concurrently():
    started, gate = [], threading.Event()

    class Slow(chat_model.Engine):
        max_batch = 64

        def __init__(self, name):
            self.name = name

        def generate(self, prompts, cancelled=None):
            started.append(self.name)
            if len(started) == 2:
                gate.set()
            assert gate.wait(5), "the replicas did not run at the same time"
            return [self.name] * len(prompts)

    r = router.ReplicaRouter([Slow("a"), Slow("b")], policy="least_loaded")
    outs = r.generate([_prompt(f"q{i}") for i in range(10)])
    assert outs.count("a") == 5 and outs.count("b") == 5
    r.close()


def test_engine_failure_propagates_and_load_is_released():
    def boom(p):
        raise RuntimeError("replica down")
    r = router.ReplicaRouter([chat_model.CallableEngine(lambda p: "ok"), chat_model.CallableEngine(boom)],
                             key_fn=lambda ps: list(range(len(ps))))
    with pytest.raises(RuntimeError, match="replica down"):
        r.generate([_prompt("a"), _prompt("b")])
    assert r.in_flight == [0, 0]
    r.close()


def test_router_behind_the_chat_model_coalesces_duplicates_per_replica():
    """B200ChatModel(engine=router): identical requests arriving together are generated once (single flight
    in the batching frontend), and the router then sees only distinct prompts."""
    calls = []

    def fn(p):
        calls.append(p)
        return "kubectl get pods"
    r = router.ReplicaRouter([chat_model.CallableEngine(fn), chat_model.CallableEngine(fn)], key_fn=_crc_keys)
    llm = chat_model.B200ChatModel(engine=r, max_wait_ms=20)
    msgs = llm.batch([_prompt("list pods")] * 6 + [_prompt("list nodes")] * 2)
    assert [m.content for m in msgs] == ["kubectl get pods"] * 8
    assert len(calls) == 2
    llm.close()


def test_bad_arguments():
    with pytest.raises(ValueError):
        router.ReplicaRouter([])
    with pytest.raises(ValueError):
        router.ReplicaRouter([chat_model.CallableEngine(lambda p: p)], policy="random")


def test_default_engine_factory_builds_a_router_for_several_replicas(monkeypatch):
    """KX_REPLICAS / KX_DEVICE steer what the unmodified app.py gets behind ChatOpenAI(...)."""
    from kubectl_b200 import engine
    made = []

    class Stub(chat_model.Engine):
        def __init__(self, arch, **kw):
            made.append(kw["device"])

    monkeypatch.setattr(engine, "B200Engine", Stub)
    monkeypatch.setenv("KX_DEVICE", "2")
    monkeypatch.delenv("KX_REPLICAS", raising=False)
    assert isinstance(engine.default_engine_factory("qwen2.5-0.5b"), Stub) and made == [2]
    monkeypatch.setenv("KX_REPLICAS", "3")
    r = engine.default_engine_factory("qwen2.5-0.5b")
    assert isinstance(r, router.ReplicaRouter) and made == [2, 2, 3, 4] and len(r.engines) == 3
    r.close()
