"""Byte-level tokenizer: the batch functions on the request path equal the scalar definitions."""
import random

import numpy as np

import kubectl_b200  # noqa: F401
from kubectl_b200 import synth, tokenizer


def test_batch_encoders_equal_the_scalar_ones():
    rng = random.Random(1)
    texts = ["", "a", "é€\U0001f600 x" * 3, "q" * 300] + ["You are... " * rng.randrange(0, 40) + q for q in synth.queries(200, seed=2)]
    for length in (1, 2, 16, 256):
        want = np.asarray([tokenizer.encode_fixed(t, length) for t in texts], np.int32)
        assert np.array_equal(want, tokenizer.encode_fixed_batch(texts, length))
        want = np.asarray([tokenizer.encode_suffix(t.encode(), length) for t in texts], np.int32)
        assert np.array_equal(want, tokenizer.encode_suffix_batch([t.encode() for t in texts], length))


def test_batch_decoder_equals_the_scalar_one():
    ids = np.random.default_rng(0).integers(0, 1024, (300, 40)).astype(np.int32)
    ids[::7, 5] = tokenizer.EOS
    ids[::5, 2] = tokenizer.BOS
    ids[3, 0] = tokenizer.EOS
    assert [tokenizer.decode(r) for r in ids] == tokenizer.decode_batch(ids)
    assert tokenizer.decode_batch(np.zeros((0, 4), np.int32)) == []
