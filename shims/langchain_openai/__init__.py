"""Zero-diff install point: put this directory ahead of site-packages on PYTHONPATH
and the unmodified reference (`from langchain_openai import ChatOpenAI`, app.py:11)
constructs the local B200 generator instead of the OpenAI client."""
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _root not in sys.path:
    sys.path.insert(0, _root)
import kubectl_b200  # noqa: E402,F401
from kubectl_b200.chat_model import B200ChatModel as ChatOpenAI  # noqa: E402,F401
