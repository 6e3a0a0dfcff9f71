"""Generation engine behind ``B200ChatModel`` (placeholder until generator lands)."""
from __future__ import annotations


def default_engine_factory(arch: str):
    raise RuntimeError("the CUDA generator engine is not available in this build")
