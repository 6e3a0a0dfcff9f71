"""Router in front of the per-GPU generator replicas.

Requests are independent, so the box runs one engine per GPU and no collective on the request
path; what is left to decide is which replica a prompt goes to.  Two policies:

* ``affinity`` (default): replica = key(prompt) mod N, where the key is the 64-bit cache-key hash
  of the *sanitised* request text -- the same SipHash-1-3 the reference's ``TTLCache`` lookup
  computes (app.py:305,312), produced for the whole batch by one k9 launch
  (``strings.sanitize_queries``).  Identical requests, and requests that differ only in the
  whitespace ``sanitize_query`` removes (app.py:60-68), therefore always meet in the same replica,
  where the batching frontend's single flight generates them once.
* ``least_loaded``: the replica with the fewest prompts in flight (ties: lowest index).

``ReplicaRouter`` is itself an ``Engine`` (chat_model.Engine), so ``B200ChatModel(engine=router)``
serves a whole box behind the reference's ``llm`` slot.  Sub-batches run concurrently, one host
thread per replica (the engines release the GIL inside the C ABI call).
"""
from __future__ import annotations

import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional, Sequence

from .chat_model import Engine
from .synth import query_of_prompt


def k9_keys(prompts: Sequence[str]) -> List[int]:
    """Cache-key hashes of the requests inside rendered prompts, from the GPU string kernel."""
    from . import strings                      # needs the CUDA library: there is no CPU fallback
    _, hashes = strings.sanitize_queries([query_of_prompt(p) for p in prompts])
    return [int(h) for h in hashes]


class ReplicaRouter(Engine):
    def __init__(self, engines: Sequence[Engine], policy: str = "affinity",
                 key_fn: Optional[Callable[[Sequence[str]], Sequence[int]]] = None):
        if not engines:
            raise ValueError("ReplicaRouter needs at least one engine")
        if policy not in ("affinity", "least_loaded"):
            raise ValueError(f"unknown routing policy {policy!r}")
        self.engines, self.policy = list(engines), policy
        self.key_fn = key_fn or k9_keys
        self.max_batch = sum(getattr(e, "max_batch", 64) for e in self.engines)
        self.in_flight = [0] * len(self.engines)
        self.routed = [0] * len(self.engines)      # prompts sent to each replica so far
        self._lock = threading.Lock()
        self._pool = ThreadPoolExecutor(max_workers=len(self.engines), thread_name_prefix="kx-replica")

    def assign(self, prompts: Sequence[str]) -> List[int]:
        n = len(self.engines)
        if n == 1:
            return [0] * len(prompts)
        if self.policy == "affinity":
            return [int(k) % n for k in self.key_fn(prompts)]
        with self._lock:
            load = list(self.in_flight)
        out = []
        for _ in prompts:
            r = min(range(n), key=lambda i: (load[i], i))
            load[r] += 1
            out.append(r)
        return out

    def generate(self, prompts: Sequence[str], cancelled=None) -> List[str]:
        where = self.assign(prompts)
        groups = [[i for i, r in enumerate(where) if r == rep] for rep in range(len(self.engines))]
        with self._lock:
            for rep, idx in enumerate(groups):
                self.in_flight[rep] += len(idx)
                self.routed[rep] += len(idx)

        def run(rep: int) -> List[str]:
            idx = groups[rep]
            try:
                eng, outs = self.engines[rep], []
                step = max(1, getattr(eng, "max_batch", len(idx)))
                for s in range(0, len(idx), step):            # a replica never sees more than its max_batch at once
                    part = idx[s:s + step]
                    outs += eng.generate([prompts[i] for i in part],
                                         None if cancelled is None else [cancelled[i] for i in part])
                if len(outs) != len(idx):
                    raise RuntimeError(f"replica {rep} returned {len(outs)} completions for {len(idx)} prompts")
                return outs
            finally:
                with self._lock:
                    self.in_flight[rep] -= len(idx)

        futures = {rep: self._pool.submit(run, rep) for rep, idx in enumerate(groups) if idx}
        result: List[Optional[str]] = [None] * len(prompts)
        for rep, fut in futures.items():
            for i, text in zip(groups[rep], fut.result()):    # an engine failure propagates (-> HTTP 500, app.py:195-197)
                result[i] = text
        return result  # type: ignore[return-value]

    def close(self) -> None:
        self._pool.shutdown(wait=True)
        for e in self.engines:
            e.close()
