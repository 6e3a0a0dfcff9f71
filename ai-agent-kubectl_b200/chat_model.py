"""Drop-in occupant of the reference's ``llm`` slot.

The reference builds ``chain = template | llm | KubectlOutputParser()`` with
``llm = ChatOpenAI(temperature=0, api_key=..., model=..., request_timeout=...,
[base_url=...])`` (app.py:108-118) and awaits ``chain.ainvoke({"query": q})`` under
``asyncio.wait_for`` (app.py:183-186).  ``B200ChatModel`` takes exactly those
keyword arguments and speaks the LangChain Runnable protocol
(``invoke / ainvoke / batch / abatch / |``), so the unmodified ``app.py`` can use it
(INTEGRATION.md).  ``api_key`` and ``base_url`` are accepted and ignored; ``model``
selects the architecture; ``request_timeout`` is a per-request deadline.

Many ``ainvoke`` coroutines run on one event loop (one per in-flight HTTP request,
app.py:299).  They are batched: each call enqueues (prompt, future) and a single
engine thread drains the queue into ``engine.generate(list_of_prompts)`` and
resolves the futures with ``loop.call_soon_threadsafe``.  When ``wait_for`` times
out it cancels the awaiting task (-> HTTP 504, app.py:189-191); the request is then
marked cancelled, dropped before scheduling, and the engine is told so it can free
that request's KV pages.

Error contract (app.py:189-197): ``ValueError`` is reserved for the parser (-> 422),
so engine failures surface as ``RuntimeError`` (-> 500 with str(e) in the detail).
"""
from __future__ import annotations

import asyncio
import logging
import queue
import threading
import time
from concurrent.futures import Future
from typing import Any, Callable, List, Optional, Sequence

logger = logging.getLogger(__name__)

try:  # real LangChain, if the deployment has it (it is not in this image)
    from langchain_core.runnables import Runnable as _LCRunnable  # type: ignore
    from langchain_core.messages import AIMessage as _AIMessage  # type: ignore
    _HAVE_LANGCHAIN = True
except Exception:  # noqa: BLE001
    _LCRunnable = object
    _AIMessage = None
    _HAVE_LANGCHAIN = False


class Message:
    """Minimal chat message (what ``StrOutputParser`` reads: ``.content``)."""

    def __init__(self, content: str, response_metadata: Optional[dict] = None):
        self.content = content
        self.type = "ai"
        self.response_metadata = response_metadata or {}

    def __repr__(self):
        return f"Message(content={self.content!r})"


def _make_message(text: str, meta: dict):
    if _HAVE_LANGCHAIN:
        return _AIMessage(content=text, response_metadata=meta)
    return Message(text, meta)


def prompt_to_text(value: Any) -> str:
    """Accepts what a Runnable in the ``llm`` position can receive: a prompt value
    (``to_string()``), a plain ``str``, or a list of messages."""
    if isinstance(value, str):
        return value
    if hasattr(value, "to_string"):
        return value.to_string()
    if isinstance(value, (list, tuple)):
        return "\n".join(prompt_to_text(getattr(m, "content", m)) for m in value)
    if hasattr(value, "content"):
        return str(value.content)
    return str(value)


class _Sequence:
    """``a | b`` when LangChain's own RunnableSequence is not available."""

    def __init__(self, *steps):
        self.steps = []
        for s in steps:
            self.steps.extend(s.steps if isinstance(s, _Sequence) else [s])

    def __or__(self, other):
        return _Sequence(self, other)

    def __ror__(self, other):
        return _Sequence(other, self)

    def invoke(self, value, config=None, **kw):
        for s in self.steps:
            value = s.invoke(value, config) if hasattr(s, "invoke") else s(value)
        return value

    async def ainvoke(self, value, config=None, **kw):
        for s in self.steps:
            if hasattr(s, "ainvoke"):
                value = await s.ainvoke(value, config)
            elif hasattr(s, "invoke"):
                value = s.invoke(value, config)
            else:
                value = s(value)
        return value

    def batch(self, values, config=None, **kw):
        values = list(values)
        for s in self.steps:
            values = s.batch(values, config) if hasattr(s, "batch") else [s.invoke(v, config) for v in values]
        return values

    async def abatch(self, values, config=None, **kw):
        return list(await asyncio.gather(*(self.ainvoke(v, config) for v in values)))


class Engine:
    """What the chat model drives.  ``generate`` maps prompts to completions for one
    batch; ``cancel`` is advisory (a request that no longer has a waiter)."""

    max_batch = 64

    def generate(self, prompts: Sequence[str], cancelled: Optional[Sequence[Callable[[], bool]]] = None) -> List[str]:
        raise NotImplementedError

    def close(self) -> None:
        pass


class CallableEngine(Engine):
    """Engine from a plain function prompt -> completion (stub LLM of config 1).  The function may return an
    exception instance to fail that one request."""

    def __init__(self, fn: Callable[[str], str], max_batch: int = 64):
        self.fn, self.max_batch, self.batches = fn, max_batch, []

    def generate(self, prompts, cancelled=None):
        self.batches.append(len(prompts))
        return [self.fn(p) for p in prompts]


class _Request:
    __slots__ = ("prompt", "future", "loop", "afuture", "deadline", "t_enqueue")

    def __init__(self, prompt, deadline):
        self.prompt, self.deadline = prompt, deadline
        self.future: Future = Future()
        self.loop = None
        self.afuture = None
        self.t_enqueue = time.monotonic()

    def cancelled(self) -> bool:
        if self.afuture is not None and self.afuture.cancelled():
            return True
        if self.future.cancelled():
            return True
        return self.deadline is not None and time.monotonic() > self.deadline


class BatchingFrontend:
    """Single consumer thread that turns concurrent requests into engine batches."""

    def __init__(self, engine: Engine, max_wait_ms: float = 2.0):
        self.engine, self.max_wait = engine, max_wait_ms / 1e3
        self.q: "queue.Queue[Optional[_Request]]" = queue.Queue()
        self.dropped = 0
        self.coalesced = 0
        self._thr = threading.Thread(target=self._run, name="kx-engine", daemon=True)
        self._thr.start()

    def submit(self, req: _Request) -> None:
        self.q.put(req)

    def stop(self):
        """Ends the consumer thread; the engine stays usable (close() also closes it)."""
        self.q.put(None)
        self._thr.join(timeout=30)

    def close(self):
        self.stop()
        self.engine.close()

    @staticmethod
    def _group_cancelled(live, prompt):
        group = [r for r in live if r.prompt == prompt]
        return lambda: all(r.cancelled() for r in group)     # only drop the work if every waiter is gone

    def _resolve(self, req: _Request, result=None, exc: Optional[BaseException] = None):
        def set_async():
            if req.afuture.done():
                return
            if exc is not None:
                req.afuture.set_exception(exc)
            else:
                req.afuture.set_result(result)

        if req.afuture is not None:
            req.loop.call_soon_threadsafe(set_async)
        elif not req.future.done():
            if exc is not None:
                req.future.set_exception(exc)
            else:
                req.future.set_result(result)

    def _run(self):
        while True:
            first = self.q.get()
            if first is None:
                return
            batch = [first]
            t_end = time.monotonic() + self.max_wait
            while len(batch) < self.engine.max_batch:
                try:
                    nxt = self.q.get(timeout=max(0.0, t_end - time.monotonic()))
                except queue.Empty:
                    break
                if nxt is None:
                    self.q.put(None)
                    break
                batch.append(nxt)
            live = []
            for r in batch:
                if r.cancelled():
                    self.dropped += 1
                    self._resolve(r, exc=TimeoutError("request cancelled before it was scheduled"))
                else:
                    live.append(r)
            if not live:
                continue
            # Single flight inside the batch: identical prompts (the reference generates twice when two
            # identical queries miss the cache concurrently, app.py:312-322) are generated once and the
            # completion is fanned out.  Greedy decoding makes the duplicates' outputs identical anyway.
            first, uniq = {}, []
            for r in live:
                if r.prompt not in first:
                    first[r.prompt] = len(uniq)
                    uniq.append(r)
            self.coalesced += len(live) - len(uniq)
            try:
                outs = self.engine.generate([r.prompt for r in uniq],
                                            [self._group_cancelled(live, r.prompt) for r in uniq])
                if len(outs) != len(uniq):
                    raise RuntimeError(f"engine returned {len(outs)} completions for {len(uniq)} prompts")
                for r in live:
                    o = outs[first[r.prompt]]
                    if isinstance(o, BaseException):       # this request alone failed (e.g. its prompt does not fit): a 500 for it only
                        self._resolve(r, exc=o if isinstance(o, RuntimeError) else RuntimeError(f"{type(o).__name__}: {o}"))
                    else:
                        self._resolve(r, result=o)
            except BaseException as e:  # noqa: BLE001 - surfaced to every waiter as a 500
                err = e if isinstance(e, RuntimeError) else RuntimeError(f"{type(e).__name__}: {e}")
                for r in live:
                    self._resolve(r, exc=err)


class SlotEngine(Engine):
    """An engine whose ``max_batch`` sequence slots are admitted, stepped and released one by one
    (``kx_engine_session_*``), which is what ``ContinuousFrontend`` drives.  All methods are called from the
    front end's single thread."""

    continuous = True
    max_new = 32

    def slots_open(self) -> None:
        raise NotImplementedError

    def slots_prepare(self, prompts: Sequence[str], idle: bool) -> List:
        """Per prompt: its token row (numpy int32 [S]), an exception instance (that request fails), or ``None``
        (cannot join the running sequences -- e.g. it does not share their cached prefix -- ask again once idle)."""
        raise NotImplementedError

    def slots_admit(self, slot0: int, rows) -> None:
        raise NotImplementedError

    def slots_step(self, n_steps: int, n_slots: int) -> None:
        raise NotImplementedError

    def slots_poll(self):
        """(tokens produced per slot, finished flag per slot) after the enqueued steps."""
        raise NotImplementedError

    def slots_read(self, slot0: int, counts: Sequence[int]) -> List[str]:
        """Completions of slots slot0 .. slot0+len(counts)-1, which have produced counts[j] tokens."""
        raise NotImplementedError

    def slots_release(self, slot0: int, n: int) -> None:
        raise NotImplementedError


def _runs(slots: Sequence[int]):
    """Sorted slot numbers -> [(first, count)] of consecutive runs."""
    out = []
    for s in sorted(slots):
        if out and out[-1][0] + out[-1][1] == s:
            out[-1][1] += 1
        else:
            out.append([s, 1])
    return [(a, n) for a, n in out]


class ContinuousFrontend(BatchingFrontend):
    """Continuous batching: requests join the running sequences at the next step boundary instead of waiting for a
    static batch to end, a finished request is answered (and its slot reused) as soon as its own sequence ends, and a
    request whose waiter has gone (``asyncio.wait_for`` timed out, app.py:183-191) gives its slot back mid-sequence.

    One loop iteration = take what arrived -> drop what was cancelled -> admit into free slots (one prefill per run of
    consecutive slots) -> ``steps_per_poll`` decode steps for every live slot -> answer the slots that ended.
    Identical prompts share one slot, also with a sequence that is already running (single flight)."""

    def __init__(self, engine: SlotEngine, max_wait_ms: float = 2.0, steps_per_poll: int = 4):
        self.steps_per_poll = max(1, int(steps_per_poll))
        self.cancelled_in_flight = 0
        self.admissions = 0
        self.ttft: List[float] = []          # seconds from submit to the first token, last 4096 requests
        self.latency: List[float] = []       # seconds from submit to the answer
        super().__init__(engine, max_wait_ms=max_wait_ms)

    def _note(self, series: List[float], value: float):
        series.append(value)
        if len(series) > 4096:
            del series[:2048]

    def _run(self):
        eng = self.engine
        n_slots_max = eng.max_batch
        free = set(range(n_slots_max))
        groups = {}                          # slot -> requests waiting for that sequence
        prompt_of, slot_of = {}, {}          # slot -> prompt, prompt -> slot
        waiting: List[_Request] = []
        opened = False
        closing = False

        def fail_all(err):
            for r in waiting:
                self._resolve(r, exc=err)
            for g in groups.values():
                for r in g:
                    self._resolve(r, exc=err)

        def free_slot(slot):
            slot_of.pop(prompt_of.pop(slot), None)
            groups.pop(slot, None)
            free.add(slot)

        while True:
            # ---- 1. intake: block only when there is nothing to do ----
            if not groups and not waiting:
                if closing:
                    return
                first = self.q.get()
                if first is None:
                    return
                waiting.append(first)
                t_end = time.monotonic() + self.max_wait        # idle: let a burst form one prefill
                while len(waiting) < n_slots_max:
                    try:
                        nxt = self.q.get(timeout=max(0.0, t_end - time.monotonic()))
                    except queue.Empty:
                        break
                    if nxt is None:
                        closing = True
                        break
                    waiting.append(nxt)
            else:
                while True:
                    try:
                        nxt = self.q.get_nowait()
                    except queue.Empty:
                        break
                    if nxt is None:
                        closing = True
                        break
                    waiting.append(nxt)
            try:
                if not opened:
                    eng.slots_open()
                    opened = True
                # ---- 2. requests nobody waits for any more ----
                keep = []
                for r in waiting:
                    if r.cancelled():
                        self.dropped += 1
                        self._resolve(r, exc=TimeoutError("request cancelled before it was scheduled"))
                    else:
                        keep.append(r)
                waiting = keep
                gone = [s for s, g in groups.items() if all(r.cancelled() for r in g)]
                for a, n in _runs(gone):
                    eng.slots_release(a, n)
                for s in gone:
                    for r in groups[s]:
                        self._resolve(r, exc=TimeoutError("request cancelled while it was decoding"))
                    self.cancelled_in_flight += 1
                    free_slot(s)
                # ---- 3. admission ----
                if waiting and free:
                    idle = not groups
                    take, index, rest = [], {}, []
                    for r in waiting:
                        if r.prompt in slot_of:                 # the same prompt is already decoding: wait for that one
                            groups[slot_of[r.prompt]].append(r)
                            self.coalesced += 1
                        elif r.prompt in index:
                            take[index[r.prompt]].append(r)
                        elif len(take) < len(free):
                            index[r.prompt] = len(take)
                            take.append([r])
                        else:
                            rest.append(r)
                    rows = eng.slots_prepare([g[0].prompt for g in take], idle) if take else []
                    ready, held = [], []
                    for g, row in zip(take, rows):
                        if isinstance(row, BaseException):
                            err = row if isinstance(row, RuntimeError) else RuntimeError(f"{type(row).__name__}: {row}")
                            for r in g:
                                self._resolve(r, exc=err)
                        elif row is None:
                            held.extend(g)
                        else:
                            ready.append((g, row))
                    # a held request needs the engine idle: nothing new is admitted ahead of it, the live slots drain
                    waiting = held + rest
                    if ready and not (held and not idle):
                        slots = sorted(free)[:len(ready)]
                        k = 0
                        for a, n in _runs(slots):
                            eng.slots_admit(a, [row for _, row in ready[k:k + n]])
                            self.admissions += 1
                            now = time.monotonic()
                            for j in range(n):
                                g = ready[k + j][0]
                                self.coalesced += len(g) - 1
                                groups[a + j] = g
                                prompt_of[a + j] = g[0].prompt
                                slot_of[g[0].prompt] = a + j
                                free.discard(a + j)
                                for r in g:
                                    self._note(self.ttft, now - r.t_enqueue)
                            k += n
                    elif ready:
                        waiting = [r for g, _ in ready for r in g] + waiting
                # ---- 4. decode steps for every live slot, answer the ones that ended ----
                if groups:
                    eng.slots_step(self.steps_per_poll, max(groups) + 1)
                    cnt, fin = eng.slots_poll()
                    ended = [s for s in groups if fin[s]]
                    for a, n in _runs(ended):
                        texts = eng.slots_read(a, [int(c) for c in cnt[a:a + n]])
                        eng.slots_release(a, n)
                        now = time.monotonic()
                        for j in range(n):
                            for r in groups[a + j]:
                                self._note(self.latency, now - r.t_enqueue)
                                self._resolve(r, result=texts[j])
                            free_slot(a + j)
            except BaseException as e:  # noqa: BLE001 - the engine failed: every waiter gets a 500, the session starts over
                err = e if isinstance(e, RuntimeError) else RuntimeError(f"{type(e).__name__}: {e}")
                logger.exception("continuous batching loop failed")
                fail_all(err)
                waiting, groups, prompt_of, slot_of = [], {}, {}, {}
                free = set(range(n_slots_max))
                opened = False
            if closing and not groups and not waiting:
                return


def make_frontend(engine: Engine, max_wait_ms: float = 2.0) -> BatchingFrontend:
    """Continuous batching when the engine has slots, static batches otherwise."""
    if getattr(engine, "continuous", False):
        return ContinuousFrontend(engine, max_wait_ms=max_wait_ms)
    return BatchingFrontend(engine, max_wait_ms=max_wait_ms)


#: names the ``model`` kwarg / OPENAI_MODEL may carry -> architecture key
ARCH_ALIASES = {
    "qwen2.5-0.5b": "qwen2.5-0.5b", "qwen/qwen2.5-0.5b": "qwen2.5-0.5b", "qwen2.5-0.5b-instruct": "qwen2.5-0.5b",
    "llama-3.2-1b": "llama-3.2-1b", "meta-llama/llama-3.2-1b": "llama-3.2-1b",
    "llama-3.1-8b": "llama-3.1-8b", "meta-llama/llama-3.1-8b": "llama-3.1-8b",
}
DEFAULT_ARCH = "qwen2.5-0.5b"


def resolve_arch(model: Optional[str]) -> str:
    key = (model or "").strip().lower()
    if key in ARCH_ALIASES:
        return ARCH_ALIASES[key]
    logger.info("model %r is not a local architecture name; using %s", model, DEFAULT_ARCH)
    return DEFAULT_ARCH


class B200ChatModel(_LCRunnable):
    """``ChatOpenAI``-compatible constructor, Runnable protocol, local generation."""

    def __init__(self, temperature: float = 0, api_key: Optional[str] = None, model: Optional[str] = None,
                 request_timeout: Optional[float] = None, base_url: Optional[str] = None, *,
                 engine: Optional[Engine] = None, engine_factory: Optional[Callable[[str], Engine]] = None,
                 max_wait_ms: float = 2.0, **ignored):
        if temperature not in (0, 0.0, None):
            raise ValueError("B200ChatModel decodes greedily; only temperature=0 is supported (app.py:109)")
        self.temperature = 0
        self.model_name = model
        self.arch = resolve_arch(model)
        self.request_timeout = request_timeout
        self.base_url, self._api_key_ignored = base_url, api_key is not None
        if base_url:
            logger.warning("B200ChatModel generates locally on the GPU: base_url=%r (and the API key) are not used", base_url)
        if engine is None:
            if engine_factory is None:
                from .engine import default_engine_factory  # the CUDA generator; raises if it cannot start
                engine_factory = default_engine_factory
            engine = engine_factory(self.arch)
        self.engine = engine
        self.frontend = make_frontend(engine, max_wait_ms=max_wait_ms)

    # ---- Runnable protocol -------------------------------------------------
    def __or__(self, other):
        if _HAVE_LANGCHAIN:
            return super().__or__(other)
        return _Sequence(self, other)

    def __ror__(self, other):
        if _HAVE_LANGCHAIN:
            return super().__ror__(other)
        return _Sequence(other, self)

    def _request(self, value) -> _Request:
        deadline = None if not self.request_timeout else time.monotonic() + float(self.request_timeout)
        return _Request(prompt_to_text(value), deadline)

    def _wrap(self, text: str):
        return _make_message(text, {"model_name": self.arch, "finish_reason": "stop"})

    def invoke(self, value, config=None, **kw):
        req = self._request(value)
        self.frontend.submit(req)
        try:
            return self._wrap(req.future.result(timeout=self.request_timeout or None))
        except BaseException:
            req.future.cancel()
            raise

    async def ainvoke(self, value, config=None, **kw):
        req = self._request(value)
        req.loop = asyncio.get_running_loop()
        req.afuture = req.loop.create_future()
        self.frontend.submit(req)
        # CancelledError from asyncio.wait_for (app.py:183-186) cancels afuture, which
        # the engine thread observes through req.cancelled().
        return self._wrap(await req.afuture)

    def batch(self, values, config=None, **kw):
        reqs = [self._request(v) for v in values]
        for r in reqs:
            self.frontend.submit(r)
        return [self._wrap(r.future.result(timeout=self.request_timeout or None)) for r in reqs]

    async def abatch(self, values, config=None, **kw):
        return list(await asyncio.gather(*(self.ainvoke(v, config) for v in values)))

    def close(self):
        self.frontend.close()

    # ---- BaseChatModel hooks (langchain_core/language_models/chat_models.py: what ChatOpenAI implements) ----
    # The service only ever calls the Runnable methods above (app.py:118, app.py:183-186); these are for callers that
    # hold the model as a chat model (callbacks, `.generate()`), and for `langchain_chat_model()` below.
    @property
    def _llm_type(self) -> str:
        return "b200-local"

    @property
    def _identifying_params(self) -> dict:
        return {"model_name": self.arch, "temperature": 0}

    @staticmethod
    def _apply_stop(text: str, stop: Optional[Sequence[str]]) -> str:
        for s in stop or ():
            k = text.find(s)
            if k >= 0:
                text = text[:k]
        return text

    def _result(self, text: str):
        msg = self._wrap(text)
        try:
            from langchain_core.outputs import ChatGeneration, ChatResult  # type: ignore
            return ChatResult(generations=[ChatGeneration(message=msg)], llm_output={"model_name": self.arch})
        except Exception:  # noqa: BLE001 - no LangChain here: same shape, plain objects
            from types import SimpleNamespace
            return SimpleNamespace(generations=[SimpleNamespace(message=msg, text=text)], llm_output={"model_name": self.arch})

    def _generate(self, messages, stop: Optional[Sequence[str]] = None, run_manager=None, **kw):
        return self._result(self._apply_stop(self.invoke(messages).content, stop))

    async def _agenerate(self, messages, stop: Optional[Sequence[str]] = None, run_manager=None, **kw):
        return self._result(self._apply_stop((await self.ainvoke(messages)).content, stop))

    def langchain_chat_model(self):
        """This model as a ``langchain_core`` ``BaseChatModel`` (needs LangChain; it is not in this image, so this
        adapter has never run against the real package -- INTEGRATION.md)."""
        from langchain_core.language_models.chat_models import BaseChatModel  # type: ignore
        inner = self

        class B200LangChainChatModel(BaseChatModel):  # type: ignore[misc]
            @property
            def _llm_type(self) -> str:
                return inner._llm_type

            @property
            def _identifying_params(self) -> dict:
                return inner._identifying_params

            def _generate(self, messages, stop=None, run_manager=None, **kw):
                return inner._generate(messages, stop=stop)

            async def _agenerate(self, messages, stop=None, run_manager=None, **kw):
                return await inner._agenerate(messages, stop=stop)

        return B200LangChainChatModel()


#: exported under the name the reference imports (app.py:11)
ChatOpenAI = B200ChatModel
