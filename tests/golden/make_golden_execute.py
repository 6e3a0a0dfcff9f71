#!/usr/bin/env python3
"""Generates tests/golden/execute.json by running the reference's own `/execute` code
(app.py:205-281, 356-389) in the build container, with the `kubectl` subprocess replaced by a
stand-in that exits 0 and prints a given stdout.  Recorded per case: the HTTP status of the
handler for the command, and the `execution_result` the reference builds from the stdout.

    python tests/golden/make_golden_execute.py
"""
import asyncio
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

STDOUTS = [
    "NAME READY STATUS RESTARTS AGE\nweb-1 1/1 Running 0 3d\nweb-2 0/1 CrashLoopBackOff 12 3d\n",
    "NAME   STATUS   ROLES    AGE   VERSION\nnode-a   Ready    <none>   10d   v1.29.1\n",
    "pod/web-1 deleted",
    "",
    "  \n",
    "NAME READY\n",
    "NAME READY\n\nweb 1/1",
    "A B C\n1 2\n3 4 5 6\n",
    "deployment.apps/web scaled\n",
    "NAME\tSTATUS\nns1\tActive\r\nns2\tTerminating",
    "Name:  x\nNamespace: y\nLabels: a=b\n",
    '{"items": []}',
    '{\n  "items": []\n}',
    "NAME Ready\nÉcole 1/1\n",
]
COMMANDS = ["kubectl get pods", "kubectl get pods; rm -rf /", " kubectl get ns ", "kubectl", "rm -rf /", "kubectl get pods -l 'a",
            "kubectl get pods | grep x", "kubectl delete pod web-1", "kubectl get $(x)", "KUBECTL get pods"]


class _Proc:
    returncode = 0

    def __init__(self, out):
        self._out = out

    async def communicate(self):
        return self._out.encode(), b""


def main():
    app = ref_loader.load_reference_app(env={"LOG_LEVEL": "ERROR"})
    from starlette.testclient import TestClient
    cases = {"stdout": [], "commands": []}
    current = {"out": ""}

    async def fake_exec(*args, **kw):
        return _Proc(current["out"])
    app.asyncio.create_subprocess_exec = fake_exec
    for out in STDOUTS:
        current["out"] = out
        res = asyncio.run(app.execute_command_async("kubectl get pods"))
        cases["stdout"].append({"stdout": out, "execution_result": res["execution_result"]})
    client = TestClient(app.app)
    current["out"] = "ok"
    for cmd in COMMANDS:
        r = client.post("/execute", json={"execute": cmd})
        cases["commands"].append({"execute": cmd, "status": r.status_code,
                                  "detail": r.json().get("detail") if r.status_code != 200 else None})
    with open(os.path.join(ROOT, "tests", "golden", "execute.json"), "w", encoding="utf-8") as f:
        json.dump(cases, f, ensure_ascii=False, indent=1)
    print(len(cases["stdout"]), "stdout cases,", len(cases["commands"]), "commands")


if __name__ == "__main__":
    main()
