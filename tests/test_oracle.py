"""Pins the CPU oracle (oracle/py_oracle.py, oracle/kx_oracle.c) to the fixtures the
verbatim reference produced (tests/golden/strings.json) and, when the reference is
present (build container only), to the reference itself."""
import os
import random
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, to_csr
from oracle import c_oracle, py_oracle as po, ref_loader


def test_spec_layer_matches_golden(golden_rows):
    for r in golden_rows:
        s = r["s"]
        assert po.spec_verdict(s) == r["reason"], repr(s)
        assert po.spec_sanitize(s) == r["san"], repr(s)
        assert po.spec_hash(r["san"]) == r["key_hash"], repr(s)
        assert po.spec_hash(s) == r["raw_hash"], repr(s)
        a, b, v = po.spec_parse(s)
        assert (s[a:b] if v == 0 else None) == r["parse"], repr(s)


def test_stdlib_layer_matches_golden(golden_rows):
    for r in golden_rows:
        s = r["s"]
        assert po.ref_verdict(s) == r["reason"], repr(s)
        assert po.ref_sanitize_query(s) == r["san"], repr(s)
        cmd, v = po.ref_parse(s)
        assert (cmd if v == 0 else None) == r["parse"], repr(s)


def test_hash_known_answers():
    # SURVEY Appendix A, hash() under PYTHONHASHSEED=0
    ka = {"list all pods": -7177872310343149011, "": 0, "a": 4644417185603328019, "é": 6047309291227476195,
          "€": -5529981157763016009, "\U0001f600": -3536540696076613844}
    for s, h in ka.items():
        assert po.spec_hash(s) == h


def test_builtin_hash_agrees_under_seed_zero():
    code = ("import sys; sys.path.insert(0, %r); from oracle import py_oracle as po; import random;"
            "r = random.Random(1); al = 'ab \\xe9\\u20ac\\U0001f600k';"
            "ss = [''.join(r.choice(al) for _ in range(r.randrange(0, 30))) for _ in range(3000)];"
            "assert po.hash_seed_is_zero(); assert all(hash(s) == po.spec_hash(s) for s in ss); print('ok')" % ROOT)
    env = dict(os.environ, PYTHONHASHSEED="0")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr


def test_c_oracle_matches_golden(golden_rows):
    strings = [r["s"] for r in golden_rows]
    data, off = to_csr(strings)
    out = c_oracle.validate_hash(data, off, flags=0, want_span=True, want_san=True, threads=2)
    outp = c_oracle.validate_hash(data, off, flags=1, want_span=True)
    for i, r in enumerate(golden_rows):
        assert out["verdict"][i] == r["reason"], repr(r["s"])
        assert out["hash"][i] == r["key_hash"], repr(r["s"])
        san = bytes(out["san"][off[i]:off[i] + out["san_len"][i]]).decode("utf-8")
        assert san == r["san"], repr(r["s"])
        raw = r["s"].encode("utf-8")
        a, b = out["span"][i]
        assert raw[a:b].decode("utf-8") == r["s"].strip()
        a, b = outp["span"][i]
        got = raw[a:b].decode("utf-8") if outp["verdict"][i] == 0 else None
        assert got == r["parse"], repr(r["s"])


def test_c_oracle_lenient_decode_matches_python_spec():
    rng = random.Random(7)
    blobs = [bytes(rng.choice([0x20, 0x27, 0x22, 0x5C, 0x60, 0x6B, 0x80, 0xA0, 0xC2, 0xE2, 0xE3, 0xF0, 0xFF, 0x28,
                               0x85, 0x9F, 0x81, 0x00]) for _ in range(rng.randrange(0, 24))) for _ in range(4000)]
    off = np.zeros(len(blobs) + 1, np.int64)
    off[1:] = np.cumsum([len(b) for b in blobs])
    data = np.frombuffer(b"".join(blobs), np.uint8)
    out = c_oracle.validate_hash(data, off, want_san=True)
    for i, b in enumerate(blobs):
        cps = po.decode_utf8_lenient(b)
        assert out["hash"][i] == po.spec_hash_codepoints(po.spec_sanitize_codepoints(cps)), b


def test_c_oracle_empty_batch_and_empty_strings():
    out = c_oracle.validate_hash(np.zeros(0, np.uint8), np.zeros(1, np.int64))
    assert out["verdict"].size == 0
    data, off = to_csr(["", "", "kubectl get pods"])
    out = c_oracle.validate_hash(data, off)
    assert list(out["verdict"]) == [1, 1, 0] and list(out["hash"][:2]) == [0, 0]


@pytest.mark.skipif(not ref_loader.reference_available(), reason="/root/reference only exists in the build container")
def test_fuzz_against_verbatim_reference():
    app = ref_loader.load_reference_app()
    rng = random.Random(99)
    al = list("kubectl get pods -n x") + list(" \t\n\\'\"`$();<>&|#") + ["\xa0", " ", "　", "é", "€", "\U0001f600"]
    strings = []
    for _ in range(20000):
        body = "".join(rng.choice(al) for _ in range(rng.randrange(0, 30)))
        strings.append(("kubectl " if rng.random() < 0.8 else "") + body)
    data, off = to_csr(strings)
    out = c_oracle.validate_hash(data, off, threads=4)
    parser = app.KubectlOutputParser()
    for i, s in enumerate(strings):
        assert (out["verdict"][i] == 0) == app.is_safe_kubectl_command(s), repr(s)
        assert po.spec_sanitize(s) == app.sanitize_query(s)
        if i % 8 == 0:
            assert out["hash"][i] == po.spec_hash(app.sanitize_query(s))
            try:
                p = parser.parse(s)
            except ValueError:
                p = None
            a, b, v = po.spec_parse(s)
            assert (s[a:b] if v == 0 else None) == p
