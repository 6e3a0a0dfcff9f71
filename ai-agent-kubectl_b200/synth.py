"""Synthetic workloads of BASELINE.json (there is no network, so no real traffic).

config 1: natural-language queries with whitespace noise and a Zipf repeat
          pattern, plus the canned completions of a stub LLM.
config 2: kubectl command strings with the mix of clean / rejected / odd cases
          SURVEY 8d row 2 specifies, as a CSR batch (uint8 data, int64 off).
All randomness is ``numpy.random.default_rng(seed)``.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

VERBS = ["get", "describe", "delete", "logs", "top", "rollout status", "scale", "label", "annotate", "edit",
         "explain", "apply -f", "patch", "cordon", "drain", "expose", "port-forward", "exec"]
RESOURCES = ["pods", "deployments", "services", "nodes", "configmaps", "secrets", "ingresses", "namespaces",
             "statefulsets", "daemonsets", "jobs", "cronjobs", "pv", "pvc", "replicasets", "events",
             "serviceaccounts", "endpoints", "hpa", "networkpolicies"]
FORMATS = ["wide", "yaml", "json", "name", "jsonpath='{.items[*].metadata.name}'",
           "custom-columns=NAME:.metadata.name,STATUS:.status.phase"]
NL_VERBS = ["list", "show", "get", "describe", "delete", "count", "watch", "find", "display", "fetch"]

_METAS = [";", "&&", "||", "`", "$", "(", ")", "<", ">"]
_UNI_WS = [" ", " ", "　", ""]
_NON_ASCII = ["é", "ü", "€", "→", "日本", "\U0001f600", "\U0001f680"]


def _clean(rng) -> str:
    s = f"kubectl {VERBS[rng.integers(len(VERBS))]} {RESOURCES[rng.integers(len(RESOURCES))]}"
    if rng.random() < 0.6:
        s += f" -n ns{rng.integers(1000)}"
    if rng.random() < 0.4:
        s += f" -l 'app=web{rng.integers(100)}'"
    if rng.random() < 0.35:
        s += f" -o {FORMATS[rng.integers(len(FORMATS))]}"
    if rng.random() < 0.1:
        s += f" --field-selector status.phase=Running --sort-by=.metadata.creationTimestamp --context cluster{rng.integers(50)}"
    return s


def command_string(rng) -> str:
    """One config-2 string (category mix from SURVEY 8d row 2)."""
    u = rng.random()
    s = _clean(rng)
    if u < 0.70:
        return s
    if u < 0.80:
        pos = int(rng.integers(8, len(s) + 1))
        return s[:pos] + _METAS[rng.integers(len(_METAS))] + s[pos:]
    if u < 0.85:
        k = int(rng.integers(6))
        body = s[len("kubectl "):]
        return ["kubectlx " + body, "Kubectl " + body, "sudo " + s, "kubectl", "kubectl\t" + body, ""][k]
    if u < 0.90:
        q = "'" if rng.random() < 0.5 else '"'
        return s.replace("'", "") + f" -l {q}tier=backend"
    if u < 0.92:
        return s.replace("'", "") + " \\"
    if u < 0.95:
        k = int(rng.integers(6))
        t = s.replace("'", "")
        return [t + " | grep Running", t + " &", t + " # note", t + "\n-o wide", t + " \\'x", t + ' "a\\" b"'][k]
    if u < 0.98:
        w = _UNI_WS[rng.integers(len(_UNI_WS))]
        return w * int(rng.integers(1, 3)) + s + w * int(rng.integers(0, 3))
    return s + " -l name=" + _NON_ASCII[rng.integers(len(_NON_ASCII))]


def command_strings(n: int, seed: int = 2) -> List[str]:
    rng = np.random.default_rng(seed)
    return [command_string(rng) for _ in range(n)]


def command_csr(n: int, seed: int = 2, pool: int = 1 << 18) -> Tuple[np.ndarray, np.ndarray]:
    """config 2 as a CSR batch.  Up to ``pool`` distinct strings are generated and
    the batch samples them uniformly with replacement (each copy still occupies its
    own bytes in the buffer).  n <= pool yields n independent strings."""
    rng = np.random.default_rng(seed)
    uniq = [command_string(rng).encode("utf-8") for _ in range(min(n, pool))]
    plen = np.fromiter((len(b) for b in uniq), np.int64, len(uniq))
    poff = np.zeros(len(uniq) + 1, np.int64)
    np.cumsum(plen, out=poff[1:])
    pdata = np.frombuffer(b"".join(uniq), np.uint8)
    if n <= pool:
        return pdata.copy(), poff
    pick = rng.integers(0, len(uniq), size=n)
    lens = plen[pick]
    off = np.zeros(n + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    data = np.empty(int(off[-1]), np.uint8)
    step = 1 << 20
    for r0 in range(0, n, step):   # ragged gather, chunked to bound the index arrays
        r1 = min(n, r0 + step)
        ln = lens[r0:r1]
        dst0 = off[r0:r1] - off[r0]
        src0 = poff[pick[r0:r1]]
        idx = np.repeat(src0 - dst0, ln) + np.arange(int(off[r1] - off[r0]), dtype=np.int64)
        data[off[r0]:off[r1]] = pdata[idx]
    return data, off


def queries(n: int = 10_000, seed: int = 1, distinct: int = 2000) -> List[str]:
    """config 1: ``n`` natural-language queries, Zipf(1.1) over ``distinct`` bases so a
    100-entry TTLCache sees hits; 20 % carry whitespace noise that sanitize_query
    must remove (so noisy and clean forms share a cache key)."""
    rng = np.random.default_rng(seed)
    bases = []
    for _ in range(distinct):
        q = f"{NL_VERBS[rng.integers(len(NL_VERBS))]} {'all' if rng.random() < 0.5 else 'the'} " \
            f"{RESOURCES[rng.integers(len(RESOURCES))]}"
        if rng.random() < 0.6:
            q += f" in namespace ns{rng.integers(1000)}"
        if rng.random() < 0.3:
            q += f" with label app=web{rng.integers(100)}"
        bases.append(q)
    ranks = np.minimum(rng.zipf(1.1, size=n), distinct) - 1
    out = []
    for r in ranks:
        q = bases[int(r)]
        if rng.random() < 0.2:
            k = int(rng.integers(5))
            q = [q.replace(" ", "  ", 1), q.replace(" ", "\t", 1), q.replace(" ", "\n", 1), "  " + q, q + " \r\n"][k]
        out.append(q)
    return out


def stub_completion(prompt: str, unsafe_rate: float = 0.05) -> str:
    """Deterministic stand-in for the LLM of config 1: a plausible command derived
    from the prompt text, deliberately unsafe for about ``unsafe_rate`` of prompts."""
    import zlib

    h = zlib.crc32(prompt.encode("utf-8"))
    res = RESOURCES[h % len(RESOURCES)]
    cmd = f"kubectl get {res} -n ns{(h >> 8) % 1000}"
    if (h >> 20) % 1000 < unsafe_rate * 1000:
        cmd += "; rm -rf /"
    return cmd


def query_of_prompt(prompt: str) -> str:
    """The request text inside a rendered prompt: what follows the last "User Request: " marker up to
    the "Kubectl Command:" cue (both markers are part of the reference's template, app.py:50-57)."""
    start = prompt.rfind("User Request: ")
    start = 0 if start < 0 else start + len("User Request: ")
    end = prompt.find("\nKubectl Command:", start)
    return prompt[start:end if end >= 0 else len(prompt)]
