"""BASELINE configs[0]: the `/kubectl-command` pipeline over 10,000 synthetic queries with a stub LLM.
The fixture (tests/golden/pipeline.json) is what the reference's own handler returned, request by
request, when executed verbatim; the pipeline mirror must reproduce it -- with the oracle's string
functions on the CPU (pipeline logic), and with the CUDA string path on a B200."""
import json
import os

import pytest

from conftest import ROOT
import kubectl_b200  # noqa: F401
from kubectl_b200 import synth
from kubectl_b200.pipeline import KubectlPipeline


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "pipeline.json")) as f:
        return json.load(f)


def _stub(prompts):
    return [synth.stub_completion(p, unsafe_rate=0.05) for p in prompts]


def _replay(pipe, golden, chunk):
    queries = synth.queries(golden["meta"]["n"], seed=golden["meta"]["query_seed"])
    got = []
    for i in range(0, len(queries), chunk):
        got.extend(pipe.handle_many(queries[i:i + chunk]))
    assert len(got) == len(golden["rows"])
    for i, (o, (status, payload, from_cache)) in enumerate(zip(got, golden["rows"])):
        assert o.status == status, (i, queries[i])
        assert (o.kubectl_command if status == 200 else o.detail) == payload, (i, queries[i])
        assert o.from_cache == from_cache, (i, queries[i])
    return got


def test_pipeline_logic_matches_the_reference_handler(golden):
    from oracle import py_oracle as po
    pipe = KubectlPipeline(_stub, render=lambda q: q,
                           sanitize_many=lambda qs: [po.ref_sanitize_query(q) for q in qs],
                           parse_many=lambda ts: tuple(zip(*[po.ref_parse(t) for t in ts])) if ts else ([], []))
    got = _replay(pipe, golden, chunk=1)
    assert sum(o.from_cache for o in got) > 5000 and any(o.status == 422 for o in got)


def test_concurrent_mode_coalesces_identical_requests():
    from oracle import py_oracle as po
    calls = []

    def gen(prompts):
        calls.append(len(prompts))
        return _stub(prompts)
    pipe = KubectlPipeline(gen, render=lambda q: q, sanitize_many=lambda qs: [po.ref_sanitize_query(q) for q in qs],
                           parse_many=lambda ts: tuple(zip(*[po.ref_parse(t) for t in ts])) if ts else ([], []))
    out = pipe.handle_concurrent(["list  pods", "list pods ", "get nodes", "list\tpods"])
    assert calls == [2] and [o.kubectl_command for o in out][0] == out[1].kubectl_command == out[3].kubectl_command
    again = pipe.handle_concurrent(["list pods"])
    assert again[0].from_cache and calls == [2]


@pytest.mark.gpu
def test_pipeline_on_the_gpu_string_path_matches_the_reference_handler(golden):
    pipe = KubectlPipeline(_stub, render=lambda q: q)          # sanitise + parse/validate on the GPU (k9)
    _replay(pipe, golden, chunk=500)
