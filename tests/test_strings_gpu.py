"""Parity of the CUDA string path (k9) with the oracle, through the C ABI."""
import random

import numpy as np
import pytest

from conftest import to_csr
from oracle import c_oracle, py_oracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ks():
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import strings
    return strings


def _all_outputs(ks, data, off, flags):
    return ks.validate_hash_host(data, off, flags, span=True, san=True)


def _same(a, b, off):
    assert np.array_equal(a["verdict"], b["verdict"])
    assert np.array_equal(a["hash"], b["hash"])
    assert np.array_equal(a["span"], b["span"])
    assert np.array_equal(a["san_len"], b["san_len"])
    for i in np.flatnonzero(a["san_len"] > 0):
        s, n = off[i], a["san_len"][i]
        assert np.array_equal(a["san"][s:s + n], b["san"][s:s + n]), i


def test_golden_fixture(ks, golden_rows):
    strings = [r["s"] for r in golden_rows]
    data, off = to_csr(strings)
    out = _all_outputs(ks, data, off, 0)
    outp = ks.validate_hash_host(data, off, 1, hash=False, span=True)
    for i, r in enumerate(golden_rows):
        raw = r["s"].encode()
        assert out["verdict"][i] == r["reason"], repr(r["s"])
        assert out["hash"][i] == r["key_hash"], repr(r["s"])
        assert bytes(out["san"][off[i]:off[i] + out["san_len"][i]]).decode() == r["san"], repr(r["s"])
        a, b = outp["span"][i]
        assert (raw[a:b].decode() if outp["verdict"][i] == 0 else None) == r["parse"], repr(r["s"])


@pytest.mark.parametrize("flags", [0, 1])
def test_synthetic_config2_sample_matches_oracle(ks, flags):
    from kubectl_b200 import synth
    data, off = synth.command_csr(200_000, seed=2)
    _same(_all_outputs(ks, data, off, flags), c_oracle.validate_hash(data, off, flags, True, True, threads=4), off)


def test_single_string_mirrors(ks):
    assert ks.sanitize_query("  list\tall\n pods \r\n") == "list all pods"
    assert ks.sanitize_query("a\xa0b c") == "a b c"
    assert ks.is_safe_kubectl_command("kubectl get pods | grep x") is True
    assert ks.is_safe_kubectl_command("kubectl get pods && x") is False
    p = ks.KubectlOutputParser()
    assert p.parse("```\nkubectl get pods\n```") == "kubectl get pods"
    assert p.parse("  kubectl get ns  \n") == "kubectl get ns"
    with pytest.raises(ValueError, match="Generated command failed safety checks: bash"):
        p.parse("```bash\nkubectl get pods\n```")
    q, h = ks.sanitize_queries(["list all pods", "", "é", "€", "\U0001f600"])
    assert list(h) == [-7177872310343149011, 0, 6047309291227476195, -5529981157763016009, -3536540696076613844]


def test_empty_and_ragged(ks):
    out = ks.validate_hash_host(np.zeros(0, np.uint8), np.zeros(1, np.int64))
    assert out["verdict"].size == 0
    data, off = to_csr(["", "", ""])
    out = ks.validate_hash_host(data, off)
    assert list(out["verdict"]) == [1, 1, 1] and list(out["hash"]) == [0, 0, 0]
    rng = random.Random(3)
    strings = ["", "kubectl get pods", " " * 300, "kubectl get " + "x" * 5000, "kubectl get 'a" + "b" * 70000,
               "　" * 100 + "kubectl get €" + " " * 50, "kubectl get pods " + "y" * 40000 + " \\"]
    strings += ["kubectl " + "z" * rng.randrange(0, 700) for _ in range(600)]   # CTA ranges beyond the staging buffer
    rng.shuffle(strings)
    data, off = to_csr(strings)
    for flags in (0, 1):
        _same(_all_outputs(ks, data, off, flags), c_oracle.validate_hash(data, off, flags, True, True), off)


def test_invalid_utf8_follows_the_documented_rule(ks):
    rng = random.Random(11)
    al = [0x20, 0x27, 0x22, 0x5C, 0x60, 0x6B, 0x80, 0xA0, 0xC2, 0xE2, 0xE3, 0xF0, 0xFF, 0x28, 0x85, 0x9F, 0x81, 0x00,
          0x40, 0xA8, 0xC1]
    blobs = [b"kubectl " * (rng.random() < 0.5) + bytes(rng.choice(al) for _ in range(rng.randrange(0, 40)))
             for _ in range(20000)]
    off = np.zeros(len(blobs) + 1, np.int64)
    off[1:] = np.cumsum([len(b) for b in blobs])
    data = np.frombuffer(b"".join(blobs), np.uint8)
    for flags in (0, 1):
        _same(_all_outputs(ks, data, off, flags), c_oracle.validate_hash(data, off, flags, True, True), off)


def test_device_entry_point_unaligned_and_offset_base(ks):
    import torch
    from kubectl_b200 import synth
    data, off = synth.command_csr(5000, seed=5)
    ref = c_oracle.validate_hash(data, off, 0, True, True)
    for shift in (0, 1, 7, 13):
        buf = torch.zeros(data.size + 64, dtype=torch.uint8, device="cuda")
        buf[shift:shift + data.size] = torch.from_numpy(data).cuda()
        # offsets that do not start at zero: base pointer is biased the other way
        bias = 1000
        d_off = torch.from_numpy(off + bias).cuda()
        view = buf[shift:]
        v = torch.empty(5000, dtype=torch.uint8, device="cuda")
        h = torch.empty(5000, dtype=torch.int64, device="cuda")
        sp = torch.empty(10000, dtype=torch.int32, device="cuda")

        class Biased:  # data_ptr() - bias, as a C caller would pass
            device = buf.device
            def numel(self): return 1
            def data_ptr(self): return view.data_ptr() - bias
        ks.validate_hash_device(Biased(), d_off, 0, verdict=v, hash=h, span=sp)
        torch.cuda.synchronize()
        assert np.array_equal(v.cpu().numpy(), ref["verdict"])
        assert np.array_equal(h.cpu().numpy(), ref["hash"])
        assert np.array_equal(sp.cpu().numpy().reshape(-1, 2), ref["span"])


def test_full_size_10m_bit_exact(ks):
    """BASELINE config 2 at full size: every verdict and hash equals the oracle's."""
    from kubectl_b200 import synth
    data, off = synth.command_csr(10_000_000, seed=2)
    out = ks.validate_hash_host(data, off)
    ref = c_oracle.validate_hash(data, off, threads=16)
    assert np.array_equal(out["verdict"], ref["verdict"])
    assert np.array_equal(out["hash"], ref["hash"])
    # size-independent property: sanitising is idempotent, so hashing the sanitised
    # strings again must reproduce the same hashes
    r2 = ks.validate_hash_host(data[:off[100000]], off[:100001], san=True)
    lens = r2["san_len"].astype(np.int64)
    off2 = np.zeros(100001, np.int64)
    np.cumsum(lens, out=off2[1:])
    idx = np.repeat(off[:100000] - off2[:-1], lens) + np.arange(int(off2[-1]))
    again = ks.validate_hash_host(r2["san"][idx], off2)
    assert np.array_equal(again["hash"], r2["hash"])


def test_router_affinity_key_is_the_cache_key_hash(ks):
    """ReplicaRouter's default key: the k9 hash of the sanitised request inside the prompt, i.e. what the
    reference's TTLCache lookup hashes (app.py:305,312) -- whitespace variants of a request share a replica."""
    import kubectl_b200  # noqa: F401
    from kubectl_b200 import chat_model, router
    head = "You are a Kubernetes CLI specialist.\nUser Request: "
    variants = ["list all pods in namespace ns7", "list  all\tpods in\nnamespace ns7 ", "\r\nlist all pods in namespace ns7"]
    others = [f"describe the node n{i}" for i in range(64)]
    prompts = [head + q + "\nKubectl Command:\n" for q in variants + others]
    r = router.ReplicaRouter([chat_model.CallableEngine(lambda p: p) for _ in range(8)])
    where = r.assign(prompts)
    assert len(set(where[:3])) == 1
    keys = router.k9_keys(prompts)
    assert keys[0] == keys[1] == keys[2] == po.spec_hash(po.ref_sanitize_query(variants[0]))   # CPython's str hash, seed 0
    assert where == [k % 8 for k in keys]
    assert len(set(where[3:])) > 4                                         # and distinct requests spread out
    r.close()


def test_lone_surrogates_are_accepted_like_the_reference_does(ks):
    """Python strings may hold lone surrogates; the reference's str functions (split, strip, hash; app.py:60-88) take
    them.  They travel as 'surrogatepass' UTF-8 (ED A0..BF xx) and come back the same way."""
    for s in ["list \ud800 pods", "  kubectl get \udfff\tpods ", "\ud83d kubectl get x"]:
        assert ks.sanitize_query(s) == po.ref_sanitize_query(s)
        assert ks.is_safe_kubectl_command(s) == (po.ref_verdict(s) == 0)
    _, hashes = ks.sanitize_queries(["list \ud800 pods"])
    assert int(hashes[0]) == po.spec_hash(po.ref_sanitize_query("list \ud800 pods"))
