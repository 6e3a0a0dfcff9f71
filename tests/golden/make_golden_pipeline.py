#!/usr/bin/env python3
"""Generates tests/golden/pipeline.json: BASELINE configs[0] -- the reference's own
`/kubectl-command` handler (app.py:299-346, executed verbatim through oracle/ref_loader.py) with a
deterministic stub in the `llm` slot, replayed sequentially over 10,000 synthetic queries.
Recorded per request: HTTP status, kubectl_command, from_cache (or the error detail).

    PYTHONHASHSEED=0 python tests/golden/make_golden_pipeline.py
"""
import asyncio
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import synth  # noqa: E402

N = 10_000


def main():
    class Stub(ref_loader.FixedResponseLLM):
        # the canned completion depends on the request only (the text between the template's
        # "User Request: " and "Kubectl Command:" markers), so the fixture does not embed the template
        def __init__(self, **kw):
            super().__init__(lambda prompt: synth.stub_completion(synth.query_of_prompt(prompt), unsafe_rate=0.05))

    app = ref_loader.load_reference_app(chat_openai=Stub, module_name="kx_ref_app_golden", env={"LOG_LEVEL": "CRITICAL"})
    from fastapi import HTTPException
    queries = synth.queries(N, seed=1)
    rows = []

    async def run():
        for q in queries:
            try:
                r = await app.get_kubectl_command(app.Query(query=q), request=None, response=None)
                rows.append([200, r.kubectl_command, r.from_cache])
            except HTTPException as e:
                rows.append([e.status_code, e.detail, False])
    asyncio.run(run())
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pipeline.json")
    tmpl = app.prompt_template_str
    meta = {"generator": "tests/golden/make_golden_pipeline.py", "n": N, "query_seed": 1,
            "stub": "kubectl_b200.synth.stub_completion(synth.query_of_prompt(prompt), unsafe_rate=0.05)",
            "template_sha256": __import__("hashlib").sha256(tmpl.encode()).hexdigest(),
            "template_len": len(tmpl), "template_prefix_len": tmpl.index("{query}")}
    with open(path, "w") as f:
        json.dump({"meta": meta, "rows": rows}, f, separators=(",", ":"))
    hits = sum(1 for r in rows if r[0] == 200 and r[2])
    print(f"wrote {len(rows)} rows ({hits} cache hits, {sum(r[0] == 422 for r in rows)} x 422) to {path}")


if __name__ == "__main__":
    main()
