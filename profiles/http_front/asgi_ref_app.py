"""ASGI entry for `uvicorn profiles.http_front.asgi_ref_app:app --workers N` (build container only): the reference's
unmodified app.py (oracle/ref_loader.py supplies stand-ins for the third-party imports that are not installed) with
B200ChatModel in its `llm` slot, forwarding to the engine process named by KX_ENGINE_SOCKET."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import kubectl_b200  # noqa: E402,F401
from kubectl_b200.chat_model import B200ChatModel  # noqa: E402
from kubectl_b200.engine_server import RemoteEngine  # noqa: E402
from oracle import ref_loader  # noqa: E402


def _llm(**kw):
    return B200ChatModel(engine=RemoteEngine(os.environ["KX_ENGINE_SOCKET"]), max_wait_ms=2.0, **kw)


_mod = ref_loader.load_reference_app(chat_openai=_llm, env={"LOG_LEVEL": "ERROR", "RATE_LIMIT": "100000000/second",
                                                            "CACHE_MAXSIZE": "1"})
app = _mod.app
