"""Engine process shared by several HTTP workers: batches span the workers' connections, one failed
request stays one failed request, and a worker's cancelled requests are not forwarded.  Stub engine, CPU only."""
import os
import tempfile
import threading
import time

import kubectl_b200  # noqa: F401
from kubectl_b200.chat_model import B200ChatModel, CallableEngine
from kubectl_b200.engine_server import EngineServer, RemoteEngine


def _server(fn, **kw):
    path = os.path.join(tempfile.mkdtemp(), "kx.sock")
    eng = CallableEngine(fn)
    return EngineServer(eng, path, **kw), eng, path


def test_batches_span_worker_connections():
    def fn(p):
        time.sleep(0.001)
        return OverflowError("too long") if "LONG" in p else "kubectl get " + p
    srv, eng, path = _server(fn, max_wait_ms=100)
    workers = [RemoteEngine(path) for _ in range(3)]
    outs = [None] * 3

    def run(i):
        outs[i] = workers[i].generate([f"w{i}-a", f"w{i}-LONG", f"w{i}-c"], [None, None, (lambda: True) if i == 0 else None])
    ts = [threading.Thread(target=run, args=(i,)) for i in range(3)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    for i in range(3):
        assert outs[i][0] == f"kubectl get w{i}-a"
        assert isinstance(outs[i][1], RuntimeError) and "too long" in str(outs[i][1])
        assert outs[i][2] == ("" if i == 0 else f"kubectl get w{i}-c")          # worker 0's cancelled request never left it
    assert srv.clients == 3 and max(eng.batches) >= 5                         # one engine batch held requests of several workers
    assert sum(eng.batches) == 8                                               # 9 prompts, one cancelled before forwarding
    for w in workers:
        w.close()
    srv.close()
    assert not os.path.exists(path)


def test_chat_model_over_a_remote_engine_and_single_flight_across_workers():
    calls = []

    def fn(p):
        calls.append(p)
        return "kubectl get pods"
    srv, eng, path = _server(fn, max_wait_ms=100)
    models = [B200ChatModel(temperature=0, api_key="x", model="m", request_timeout=30, engine=RemoteEngine(path)) for _ in range(2)]
    res = [None, None]

    def run(i):
        res[i] = [m.content for m in models[i].batch(["same prompt", f"own prompt {i}"])]
    ts = [threading.Thread(target=run, args=(i,)) for i in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert res[0] == res[1] == ["kubectl get pods", "kubectl get pods"]
    assert sorted(calls) == ["own prompt 0", "own prompt 1", "same prompt"]    # the shared prompt was generated once for both workers
    for m in models:
        m.close()
    srv.close()
