"""The `/execute` side of the reference: its call into the validator (app.py:375-379) and its parser
of `kubectl`'s table output (app.py:231-249), against outputs recorded from the verbatim reference
(tests/golden/execute.json, made by tests/golden/make_golden_execute.py)."""
import json
import os

import pytest

import kubectl_b200  # noqa: F401
from conftest import ROOT
from kubectl_b200 import execute


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "execute.json"), encoding="utf-8") as f:
        return json.load(f)


def test_table_parser_matches_the_reference(golden):
    for case in golden["stdout"]:
        assert execute.parse_kubectl_output(case["stdout"]) == case["execution_result"], case["stdout"]
    assert execute.parse_kubectl_outputs([c["stdout"] for c in golden["stdout"]]) == [c["execution_result"] for c in golden["stdout"]]


@pytest.mark.gpu
def test_execute_requests_are_checked_on_the_gpu_like_app_py_375(golden):
    cmds = [c["execute"] for c in golden["commands"]]
    got = execute.check_execute_requests(cmds)
    for c, detail in zip(golden["commands"], got):
        assert (detail is None) == (c["status"] == 200), c
        if detail is not None:
            assert c["status"] == execute.REJECT_STATUS and detail == c["detail"]
