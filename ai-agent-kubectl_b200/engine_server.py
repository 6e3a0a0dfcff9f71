"""One engine per GPU, many HTTP worker processes in front of it.

The reference serves every request on ONE asyncio loop (app.py:391-400): FastAPI routing, pydantic,
the rate limiter and the Prometheus middleware cost about 5 ms of host time per request, i.e. about
185 requests/s per process, far below what a GPU replica generates.  The reference's own remedy for
that is more uvicorn workers -- but each worker process would then construct its own ``llm``
(app.py:108-117), i.e. its own copy of the model on the GPU, and batches would never span workers.

This module splits the two: ``serve()`` runs in ONE process per GPU (or per box, behind a
``ReplicaRouter``) and owns the engine; every uvicorn worker runs the unmodified ``app.py`` with a
``RemoteEngine`` in its ``B200ChatModel`` (select it with ``KX_ENGINE_SOCKET=/path``, see
``engine.default_engine_factory``).  Requests of all workers meet in the server's batching front end,
so a batch spans workers, identical prompts are generated once across workers, and the HTTP overhead
scales with the number of worker processes instead of capping the box.

Transport: ``multiprocessing.connection`` over a Unix socket (length-prefixed pickles, authenticated
with a shared key); one connection per worker, one server thread per connection.  A message is the
list of prompts of one worker-side batch; the reply is the list of completions, where an element may
be an exception instance (that request alone failed, cf. ``chat_model.BatchingFrontend``).
"""
from __future__ import annotations

import logging
import os
import threading
from multiprocessing.connection import Client, Listener
from typing import List, Optional, Sequence

from .chat_model import Engine, _Request, make_frontend

logger = logging.getLogger(__name__)
_AUTHKEY = b"kx-engine"


class EngineServer:
    """Owns an engine and serves ``RemoteEngine`` clients on a Unix socket."""

    def __init__(self, engine: Engine, address: str, max_wait_ms: float = 2.0, authkey: bytes = _AUTHKEY):
        if os.path.exists(address):
            os.unlink(address)
        self.address = address
        self.frontend = make_frontend(engine, max_wait_ms=max_wait_ms)   # continuous batching if the engine has slots
        self.listener = Listener(address, family="AF_UNIX", authkey=authkey)
        self.clients = 0
        self._closing = False
        self._thr = threading.Thread(target=self._accept, name="kx-engine-server", daemon=True)
        self._thr.start()

    def _accept(self):
        while not self._closing:
            try:
                conn = self.listener.accept()
            except (OSError, EOFError):
                return
            self.clients += 1
            threading.Thread(target=self._serve_client, args=(conn,), daemon=True).start()

    def _serve_client(self, conn):
        try:
            while True:
                try:
                    prompts = conn.recv()
                except (EOFError, OSError):
                    return
                reqs = [_Request(p, None) for p in prompts]
                for r in reqs:
                    self.frontend.submit(r)          # batches form across all connections
                out = []
                for r in reqs:
                    try:
                        out.append(r.future.result())
                    except BaseException as e:  # noqa: BLE001 - travels to the worker, which raises it for that request
                        out.append(e if isinstance(e, RuntimeError) else RuntimeError(f"{type(e).__name__}: {e}"))
                conn.send(out)
        finally:
            conn.close()

    def close(self):
        self._closing = True
        try:
            self.listener.close()
        finally:
            self.frontend.close()
            if os.path.exists(self.address):
                os.unlink(self.address)


class RemoteEngine(Engine):
    """``chat_model.Engine`` that forwards batches to an ``EngineServer``."""

    def __init__(self, address: str, max_batch: int = 512, authkey: bytes = _AUTHKEY):
        self.address, self.max_batch, self._authkey = address, max_batch, authkey
        self._conn = None
        self._lock = threading.Lock()

    def generate(self, prompts: Sequence[str], cancelled=None) -> List:
        out: List = [None] * len(prompts)
        live = [i for i in range(len(prompts))
                if not (cancelled is not None and cancelled[i] is not None and cancelled[i]())]
        for i in range(len(prompts)):
            if i not in live:
                out[i] = ""
        if not live:
            return out
        with self._lock:
            if self._conn is None:
                self._conn = Client(self.address, family="AF_UNIX", authkey=self._authkey)
            self._conn.send([prompts[i] for i in live])
            reply = self._conn.recv()
        if len(reply) != len(live):
            raise RuntimeError(f"engine server returned {len(reply)} completions for {len(live)} prompts")
        for i, o in zip(live, reply):
            out[i] = o
        return out

    def close(self):
        with self._lock:
            if self._conn is not None:
                self._conn.close()
                self._conn = None


def main(argv: Optional[Sequence[str]] = None) -> None:
    import argparse
    import signal
    ap = argparse.ArgumentParser(description="engine process for several uvicorn workers (KX_ENGINE_SOCKET)")
    ap.add_argument("--arch", default="qwen2.5-0.5b")
    ap.add_argument("--socket", default="/tmp/kx-engine.sock")
    ap.add_argument("--max-wait-ms", type=float, default=2.0)
    args = ap.parse_args(argv)
    os.environ.pop("KX_ENGINE_SOCKET", None)            # this process builds the real engine(s): KX_DEVICE / KX_REPLICAS apply
    from .engine import default_engine_factory
    srv = EngineServer(default_engine_factory(args.arch), args.socket, max_wait_ms=args.max_wait_ms)
    logger.warning("engine server for %s listening on %s", args.arch, args.socket)
    signal.sigwait({signal.SIGINT, signal.SIGTERM})
    srv.close()


if __name__ == "__main__":
    main()
