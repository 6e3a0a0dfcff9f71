"""TEST INFRASTRUCTURE ONLY -- loads the reference's own ``app.py`` verbatim.

Nothing on the product path may import this module (see ``oracle/README.md``).

The reference (``/root/reference/app.py``) cannot be imported as shipped in this
container: ``langchain``, ``langchain_core``, ``langchain_openai`` and ``slowapi``
are neither installed nor in the offline wheelhouse.  This loader puts minimal
stand-ins for exactly the names ``app.py:11-14`` and ``app.py:17-20`` import into
``sys.modules`` (only for packages that are really missing) and then executes the
reference file unmodified through ``importlib``.  Under it ``sanitize_query``
(``app.py:60-68``), ``is_safe_kubectl_command`` (``app.py:72-88``),
``KubectlOutputParser.parse`` (``app.py:95-104``) and the ``/kubectl-command``
handler (``app.py:299-346``) are the reference's own bytes.

``/root/reference`` exists only in the build container, never on the GPU box, so
this is used (a) by ``tests/golden/make_golden.py`` to produce the committed
fixtures and (b) by CPU tests that skip when the path is absent.

The stand-ins implement the slice of the LangChain Runnable protocol the
reference touches: ``a | b`` composition, ``invoke``/``ainvoke``, a prompt value
with ``to_string()`` and an output parser that accepts a ``str`` or an object
with ``.content``.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys
import types
from typing import Any, Callable, Optional

REFERENCE_APP = os.environ.get("KX_REFERENCE_APP", "/root/reference/app.py")


def reference_available() -> bool:
    return os.path.isfile(REFERENCE_APP)


# --------------------------------------------------------------------------- #
# Stand-in Runnable protocol (langchain-core 0.3 surface used by app.py:118)
# --------------------------------------------------------------------------- #
class _Runnable:
    def __or__(self, other):
        return _Sequence(self, other)

    def __ror__(self, other):
        return _Sequence(other, self)

    def invoke(self, value, config=None, **kw):  # pragma: no cover - abstract
        raise NotImplementedError

    async def ainvoke(self, value, config=None, **kw):
        return self.invoke(value, config, **kw)

    def batch(self, values, config=None, **kw):
        return [self.invoke(v, config, **kw) for v in values]

    async def abatch(self, values, config=None, **kw):
        import asyncio

        return list(await asyncio.gather(*(self.ainvoke(v, config, **kw) for v in values)))


class _Sequence(_Runnable):
    def __init__(self, *steps):
        flat = []
        for s in steps:
            flat.extend(s.steps if isinstance(s, _Sequence) else [s])
        self.steps = flat

    def invoke(self, value, config=None, **kw):
        for s in self.steps:
            value = s.invoke(value, config)
        return value

    async def ainvoke(self, value, config=None, **kw):
        for s in self.steps:
            if hasattr(s, "ainvoke"):
                value = await s.ainvoke(value, config)
            else:
                value = s.invoke(value, config)
        return value

    def batch(self, values, config=None, **kw):
        for s in self.steps:
            if hasattr(s, "batch"):
                values = s.batch(values, config)
            else:
                values = [s.invoke(v, config) for v in values]
        return values


class _StringPromptValue:
    def __init__(self, text: str):
        self.text = text

    def to_string(self) -> str:
        return self.text

    def __str__(self) -> str:
        return self.text


class _PromptTemplate(_Runnable):
    def __init__(self, input_variables=None, template: str = "", **_):
        self.input_variables = list(input_variables or [])
        self.template = template

    def format(self, **kw) -> str:
        return self.template.format(**kw)

    def invoke(self, value, config=None, **kw):
        return _StringPromptValue(self.format(**value))


class _StrOutputParser(_Runnable):
    def parse(self, text: str) -> str:
        return text

    def invoke(self, value, config=None, **kw):
        text = value.content if hasattr(value, "content") else value
        return self.parse(text)


class _LLMChain:  # dead import at app.py:13
    def __init__(self, *a, **k):
        pass


class _OfflineChatOpenAI(_Runnable):
    """Default occupant of the ``llm`` slot: there is no network, so it raises."""

    def __init__(self, **kwargs):
        self.kwargs = kwargs

    def invoke(self, value, config=None, **kw):
        raise RuntimeError("OpenAI is unreachable offline; install a stub in the llm slot")


class FixedResponseLLM(_Runnable):
    """Equivalent of LangChain's FakeListLLM (not installed here): maps the
    rendered prompt to a canned completion through ``fn``."""

    def __init__(self, fn: Callable[[str], str]):
        self.fn = fn
        self.calls = 0

    def invoke(self, value, config=None, **kw):
        self.calls += 1
        text = value.to_string() if hasattr(value, "to_string") else str(value)
        return self.fn(text)


# --------------------------------------------------------------------------- #
# Stand-in slowapi (rate limiting is out of scope; limiter must not cap tests)
# --------------------------------------------------------------------------- #
class _Limiter:
    def __init__(self, key_func=None, default_limits=None, **_):
        self.key_func = key_func
        self.default_limits = default_limits

    def limit(self, *_a, **_k):
        def deco(fn):
            return fn

        return deco


class _RateLimitExceeded(Exception):
    pass


def _rate_limit_exceeded_handler(request, exc):  # pragma: no cover
    from starlette.responses import JSONResponse

    return JSONResponse({"error": "rate limit exceeded"}, status_code=429)


def _get_remote_address(request):  # pragma: no cover
    return getattr(getattr(request, "client", None), "host", "127.0.0.1")


class _SlowAPIMiddleware:
    def __init__(self, app, **_):
        self.app = app

    async def __call__(self, scope, receive, send):
        await self.app(scope, receive, send)


def _missing(name: str) -> bool:
    try:
        return importlib.util.find_spec(name) is None
    except (ImportError, ValueError):
        return True


def _mod(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__kx_stub__ = True
    sys.modules[name] = m
    return m


def install_stubs(chat_openai: Optional[type] = None) -> None:
    """Install stand-ins for the third-party imports of app.py that are missing.
    ``chat_openai`` replaces the class bound to ``langchain_openai.ChatOpenAI``
    (this is the zero-diff install point of the drop-in)."""
    if _missing("langchain_core") or getattr(sys.modules.get("langchain_core"), "__kx_stub__", False):
        pkg = _mod("langchain_core")
        pkg.__path__ = []
        pkg.output_parsers = _mod("langchain_core.output_parsers", StrOutputParser=_StrOutputParser)
    if _missing("langchain") or getattr(sys.modules.get("langchain"), "__kx_stub__", False):
        pkg = _mod("langchain")
        pkg.__path__ = []
        pkg.prompts = _mod("langchain.prompts", PromptTemplate=_PromptTemplate)
        pkg.chains = _mod("langchain.chains", LLMChain=_LLMChain)
    cur = sys.modules.get("langchain_openai")
    if chat_openai is not None:
        _mod("langchain_openai", ChatOpenAI=chat_openai)
    elif cur is None and _missing("langchain_openai"):
        _mod("langchain_openai", ChatOpenAI=_OfflineChatOpenAI)
    if _missing("slowapi") or getattr(sys.modules.get("slowapi"), "__kx_stub__", False):
        pkg = _mod("slowapi", Limiter=_Limiter, _rate_limit_exceeded_handler=_rate_limit_exceeded_handler)
        pkg.__path__ = []
        pkg.util = _mod("slowapi.util", get_remote_address=_get_remote_address)
        pkg.errors = _mod("slowapi.errors", RateLimitExceeded=_RateLimitExceeded)
        pkg.middleware = _mod("slowapi.middleware", SlowAPIMiddleware=_SlowAPIMiddleware)


def load_reference_app(chat_openai: Optional[type] = None, env: Optional[dict] = None,
                       module_name: str = "kx_reference_app") -> Any:
    """Execute ``/root/reference/app.py`` unmodified and return the module.

    ``env`` entries are set in ``os.environ`` before the exec because app.py reads
    its configuration at import time (``app.py:24-36``)."""
    if not reference_available():
        raise FileNotFoundError(
            f"{REFERENCE_APP} not found: the verbatim reference only exists in the build container")
    defaults = {"LOG_LEVEL": "ERROR", "OPENAI_API_KEY": "offline", "RATE_LIMIT": "1000000/second"}
    for k, v in {**defaults, **(env or {})}.items():
        os.environ[k] = str(v)
    install_stubs(chat_openai)
    spec = importlib.util.spec_from_file_location(module_name, REFERENCE_APP)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[module_name] = mod
    spec.loader.exec_module(mod)
    return mod
