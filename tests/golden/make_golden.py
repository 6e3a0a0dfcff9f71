#!/usr/bin/env python3
"""Generates tests/golden/strings.json by executing the reference's own code.

Run in the build container only (``/root/reference`` is not on the GPU box):

    PYTHONHASHSEED=0 python tests/golden/make_golden.py

For every input string the fixture records what the verbatim reference returns:
  safe      is_safe_kubectl_command(s)                      app.py:72-88
  reason    which of its three rejection branches logged    app.py:76,80,86
  san       sanitize_query(s)                               app.py:60-68
  key_hash  hash(sanitize_query(s)) -- the dict hash TTLCache uses, app.py:312
  parse     KubectlOutputParser().parse(s) or null on ValueError   app.py:95-104
Inputs: the SURVEY Appendix A table, hand-written edge cases, and a seeded fuzz
over an alphabet biased to the characters the validator cares about.
"""
import json
import logging
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

if os.environ.get("PYTHONHASHSEED") != "0":
    os.environ["PYTHONHASHSEED"] = "0"
    os.execv(sys.executable, [sys.executable] + sys.argv)

from oracle import ref_loader  # noqa: E402

CURATED = [
    'kubectl get pods', ' kubectl get pods ', 'kubectl', 'kubectl ', 'kubectl\tget pods', 'kubectlx get',
    'KUBECTL get pods', 'kubectl get pods; rm -rf /',
    "kubectl get pods -o jsonpath='{.items[*].metadata.name}'", "kubectl get pods -l 'app=web",
    'kubectl get pods -l "app=web', 'kubectl get po \\', "kubectl exec pod -- sh -c 'echo $HOME'",
    'kubectl get pods | grep x', 'kubectl get pods & ', 'kubectl get pods && x', 'kubectl logs a > f',
    'kubectl get pods\nrm -rf', "kubectl get pods #comment 'unclosed", "kubectl get pods \\'",
    "kubectl get 'a\\'", 'kubectl get "a\\"', 'kubectl get "a\\" b"', '\xa0kubectl get pods',
    'kubectl\xa0get', 'kubectl get pods ', '```kubectl get pods```', "kubectl get é'",
    '```\nkubectl get pods\n```', '```bash\nkubectl get pods\n```', '  kubectl get ns  \n',
    '', ' ', '\t\n', '```', '````', '`````', '``````', '``` ```', '```kubectl get pods``` ', '``` kubectl get x 　```',
    '```kubectl get pods``', '``kubectl get pods```', 'kubectl get pods ||true', 'kubectl get | |', 'kubectl get &|&',
    'kubectl get pods \\ ', 'kubectl get pods \\\\', 'kubectl get pods \\\\\\', 'kubectl get "\\', 'kubectl get "\\\\"',
    "kubectl get ''", 'kubectl get ""', "kubectl get '\"'", 'kubectl get "\'"', "kubectl get '\"", 'kubectl  get', 'kubectl  ',
    'kubectl x', 'kubectl x', '\x1ckubectl x\x1f', 'kubectl \x1c', 'sudo kubectl get pods', 'kubectl get \U0001f600',
    'kubectl get €', 'kubectl get é', '　kubectl get pods　', 'kubectl get pods  ', 'kubectl get pods',
    'kubectl get pods ', '  list\tall\n pods \r\n', 'a\xa0b c', 'x\x0by\x0cz\x1cw', 'list all pods', 'a',
    'é', '€', '\U0001f600', 'a \U0001f600 b', 'ÿ', 'Ā', '￿', '\U00010000', 'abcdefgh', 'abcdefg', 'abcdefghi',
    'abcdefghijklmnop', '€' * 4, '€' * 5, '\U0001f600' * 2, '\U0001f600' * 3, 'kubectl get pods -n kube-system -o wide',
    'kubectl get \x00', '\x00', 'kubectl get pods\x00\\', '​', 'kubectl get ​', '᠎ x', '﻿kubectl get pods',
]

ALPHABET = (list("kubectl get pods -n ns -l app=web -o json") * 2 + list(" \t\n\r\x0b\x0c\x1c\x1d\x1e\x1f") +
            list("\\\\''\"\"`$();<>&&||##") + ["\x85", "\xa0", " ", " ", " ", " ", " ",
                                                  " ", " ", "　", "​", "é", "€",
                                                  "\U0001f600", "\x00", "\x7f", "\xff", "Ā"])


def fuzz(rng, n):
    out = []
    for _ in range(n):
        kind = rng.random()
        body = "".join(rng.choice(ALPHABET) for _ in range(rng.randrange(0, 40)))
        if kind < 0.55:
            s = "kubectl " + body
        elif kind < 0.7:
            s = rng.choice([" ", "\n", "　", "\xa0", ""]) + "kubectl " + body + rng.choice([" ", "\t", " ", ""])
        elif kind < 0.8:
            s = "```" + rng.choice(["", "\n", " "]) + "kubectl " + body + rng.choice(["", "\n", " "]) + "```"
        else:
            s = body
        out.append(s)
    return out


def main():
    app = ref_loader.load_reference_app()
    reasons = []

    class Grab(logging.Handler):
        def emit(self, rec):
            m = rec.getMessage()
            reasons.append(1 if "does not start with" in m else 2 if "potentially unsafe" in m else
                           3 if "failed shlex parsing" in m else -1)

    app.logger.setLevel(logging.WARNING)
    app.logger.addHandler(Grab())
    app.logger.propagate = False
    parser = app.KubectlOutputParser()

    rng = random.Random(20261106)
    inputs = list(dict.fromkeys(CURATED + fuzz(rng, 4000)))
    rows = []
    for s in inputs:
        reasons.clear()
        safe = app.is_safe_kubectl_command(s)
        reason = 0 if safe else reasons[-1]
        assert (reason == 0) == safe and reason >= 0
        san = app.sanitize_query(s)
        try:
            parsed = parser.parse(s)
        except ValueError:
            parsed = None
        rows.append({"s": s, "safe": safe, "reason": reason, "san": san, "key_hash": hash(san),
                     "raw_hash": hash(s), "parse": parsed})
    meta = {"generator": "tests/golden/make_golden.py", "reference": "app.py (verbatim, via oracle/ref_loader.py)",
            "python": sys.version.split()[0], "PYTHONHASHSEED": "0", "n": len(rows)}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "strings.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump({"meta": meta, "rows": rows}, f, ensure_ascii=True, separators=(",", ":"))
    print(f"wrote {len(rows)} rows to {path}")


if __name__ == "__main__":
    main()
