"""HTTP front-end scaling, build container only (needs /root/reference): the unmodified app.py under
uvicorn with 1, 2, 4 ... worker processes, all forwarding to ONE engine process (stub engine: a fixed safe command
after 2 ms per batch, so the HTTP side is what is measured).  Every request is a distinct query (cache misses).
usage: python profiles/http_front/run.py [workers ...]"""
import http.client
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import kubectl_b200  # noqa: E402,F401
from kubectl_b200.chat_model import CallableEngine  # noqa: E402
from kubectl_b200.engine_server import EngineServer  # noqa: E402

PORT = 18765


class SlowBatch(CallableEngine):
    def generate(self, prompts, cancelled=None):
        time.sleep(0.002)
        return super().generate(prompts, cancelled)


def load(n_threads, per_thread, tag):
    done = [0] * n_threads

    def run(t):
        c = http.client.HTTPConnection("127.0.0.1", PORT)
        for i in range(per_thread):
            body = json.dumps({"query": f"list pods in namespace ns{tag}-{t}-{i}"})
            c.request("POST", "/kubectl-command", body, {"Content-Type": "application/json"})
            r = c.getresponse()
            r.read()
            done[t] += r.status == 200
        c.close()
    ts = [threading.Thread(target=run, args=(t,)) for t in range(n_threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return sum(done), time.perf_counter() - t0


def main():
    sock = os.path.join(tempfile.mkdtemp(), "kx.sock")
    eng = SlowBatch(lambda p: "kubectl get pods -n kube-system", max_batch=512)
    srv = EngineServer(eng, sock, max_wait_ms=2.0)
    env = dict(os.environ, KX_ENGINE_SOCKET=sock, PYTHONPATH=ROOT)
    for w in [int(a) for a in sys.argv[1:]] or [1, 2, 4]:
        p = subprocess.Popen([sys.executable, "-m", "uvicorn", "profiles.http_front.asgi_ref_app:app", "--port", str(PORT),
                              "--workers", str(w), "--log-level", "error"], cwd=ROOT, env=env)
        try:
            for _ in range(100):                          # wait for the workers
                try:
                    c = http.client.HTTPConnection("127.0.0.1", PORT, timeout=1)
                    c.request("GET", "/health"); c.getresponse().read(); c.close()
                    break
                except OSError:
                    time.sleep(0.2)
            time.sleep(1.0)
            load(8, 50, f"warm{w}")
            b0 = len(eng.batches)
            import multiprocessing as mp                  # several client processes: one Python process cannot offer enough load
            t0 = time.perf_counter()
            with mp.get_context("fork").Pool(4) as pool:
                parts = pool.starmap(load, [(16, 150, f"w{w}c{c}") for c in range(4)])
            ok, dt = sum(p_[0] for p_ in parts), time.perf_counter() - t0
            sizes = eng.batches[b0:]
            print(f"workers {w}: {ok} requests (all 200) in {dt:.2f} s = {ok / dt:.0f} requests/s; engine batches: {len(sizes)}, "
                  f"mean size {sum(sizes) / max(1, len(sizes)):.1f}, largest {max(sizes)}", flush=True)
        finally:
            p.terminate()
            p.wait(timeout=10)
            time.sleep(0.5)
    srv.close()


if __name__ == "__main__":
    main()
