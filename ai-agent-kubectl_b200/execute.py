"""The `/execute` side of the reference as batch utilities.

The reference's `/execute` handler (app.py:356-389) does two pieces of string work around the
`kubectl` subprocess it spawns:

  * before: ``is_safe_kubectl_command(req.execute)``, else HTTP 400 "Command failed safety
    checks" (app.py:375-379) -- the same validator the output parser uses;
  * after: the successful command's stdout is turned into ``{"type": "table", "data": [...]}``
    when it has more than one line (first line = column headers, lower-cased; every other line
    zipped with them) and into ``{"type": "raw", "data": stdout}`` otherwise (app.py:231-249).

``check_execute_requests`` runs the first for a whole batch of requests on the GPU (kernel k9);
``parse_kubectl_output`` / ``parse_kubectl_outputs`` restate the second.  Spawning ``kubectl``
itself stays the reference's code (it needs a cluster).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

from . import strings

REJECT_STATUS = 400                                   # app.py:376-379
REJECT_DETAIL = "Command failed safety checks"


def check_execute_requests(commands: Sequence[str], device: int = 0) -> List[Optional[str]]:
    """Per request: ``None`` if the command may be executed, else the detail of the 400 the
    reference answers with.  One kernel launch for the batch."""
    verdicts = strings.command_verdicts(list(commands), device=device)
    return [None if v == strings.SAFE else REJECT_DETAIL for v in verdicts]


def parse_kubectl_output(stdout: str) -> Dict[str, Any]:
    """``execution_result`` for the decoded stdout of a command that exited 0 (app.py:231-249)."""
    text = stdout.strip()
    if "\n" not in text:
        return {"type": "raw", "data": text}
    rows = text.splitlines()
    columns = [name.lower() for name in rows[0].split()]
    return {"type": "table", "data": [dict(zip(columns, row.split())) for row in rows[1:]]}


def parse_kubectl_outputs(stdouts: Sequence[str]) -> List[Dict[str, Any]]:
    return [parse_kubectl_output(s) for s in stdouts]
