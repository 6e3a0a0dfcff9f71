"""The `/kubectl-command` request path of the reference (app.py:299-346), as a library.

``KubectlPipeline.handle`` mirrors the handler step by step -- sanitise -> cache probe ->
(miss) generate -> parse + validate -> cache store -- with the same cache class and keying
(``cachetools.TTLCache(maxsize, ttl)`` keyed by the sanitised query, app.py:125,312-322), the
same outcomes (a validated command, ``from_cache``; an unsafe completion is an error and is not
cached, app.py:130-131 of SURVEY 3.1) and the reference's error texts.  The string work runs on
the GPU: ``handle_many`` sanitises every query of a batch in one kernel launch (cache keys) and
validates all generated completions in another, while the cache is still walked request by
request so that a replay gives exactly what the reference gives sequentially.

Nothing here talks HTTP; FastAPI, auth, rate limiting and metrics stay the reference's own code.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

from cachetools import TTLCache

#: app.py:50-57 -- the service renders its prompt with the reference's own template at run time;
#: a pipeline built outside the service is given the renderer by its caller.
PromptFn = Callable[[str], str]


@dataclass
class Outcome:
    status: int                      # 200, or 422 when the completion fails the safety checks (app.py:192-194)
    kubectl_command: Optional[str]
    from_cache: bool
    detail: Optional[str] = None


class KubectlPipeline:
    def __init__(self, generate: Callable[[Sequence[str]], List[str]], render: PromptFn,
                 sanitize_many: Optional[Callable[[Sequence[str]], List[str]]] = None,
                 parse_many: Optional[Callable[[Sequence[str]], tuple]] = None,
                 cache_maxsize: int = 100, cache_ttl: int = 300):
        if sanitize_many is None or parse_many is None:
            from . import strings                       # the CUDA string path; no CPU fallback
            sanitize_many = sanitize_many or (lambda qs: strings.sanitize_queries(qs)[0])
            parse_many = parse_many or strings.parse_outputs
        self.generate, self.render = generate, render
        self.sanitize_many, self.parse_many = sanitize_many, parse_many
        self.cache = TTLCache(maxsize=cache_maxsize, ttl=cache_ttl)        # app.py:125

    def _finish(self, key: str, text: str, command: str, verdict: int) -> Outcome:
        if verdict != 0:                                                    # app.py:102-103, 192-194
            return Outcome(422, None, False,
                           f"LLM generated unsafe command: Generated command failed safety checks: {command}")
        self.cache[key] = command                                           # app.py:322
        return Outcome(200, command, False)

    def handle(self, query: str) -> Outcome:
        return self.handle_many([query])[0]

    def handle_many(self, queries: Sequence[str]) -> List[Outcome]:
        """Requests in arrival order.  Keys for the whole batch come from one GPU launch; misses are
        generated and validated as they are met, so the cache evolves exactly as in a sequential run."""
        keys = self.sanitize_many(list(queries))                            # app.py:305
        out: List[Outcome] = []
        for key in keys:
            hit = self.cache.get(key)                                       # app.py:312
            if hit is not None:
                out.append(Outcome(200, hit, True))
                continue
            text = self.generate([self.render(key)])[0]                     # app.py:320 -> chain (app.py:118)
            cmds, verdicts = self.parse_many([text])                        # KubectlOutputParser.parse, app.py:95-104
            out.append(self._finish(key, text, cmds[0], int(verdicts[0])))
        return out

    def handle_concurrent(self, queries: Sequence[str]) -> List[Outcome]:
        """All requests arrive at once (the reference under load: every identical query that misses
        generates, app.py:312-322 has an await between probe and store).  Misses are generated as ONE
        batch and validated in ONE launch; identical keys are generated once (single flight)."""
        keys = self.sanitize_many(list(queries))
        hits = [self.cache.get(k) for k in keys]
        todo = list(dict.fromkeys(k for k, h in zip(keys, hits) if h is None))
        texts = self.generate([self.render(k) for k in todo]) if todo else []
        cmds, verdicts = self.parse_many(texts) if todo else ([], [])
        done = {k: self._finish(k, t, c, int(v)) for k, t, c, v in zip(todo, texts, cmds, verdicts)}
        return [Outcome(200, h, True) if h is not None else done[k] for k, h in zip(keys, hits)]
