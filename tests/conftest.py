import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def random_ct_words(rng, batch, W, n, k):
    """uint64[batch][2][W][n]: uniform two's-complement coefficients in [-2^(k-1), 2^(k-1))."""
    w = rng.integers(0, 1 << 64, size=(batch, 2, W, n), dtype=np.uint64)
    top_bits = k - 64 * (W - 1)
    assert 1 <= top_bits <= 64
    if top_bits < 64:
        t = w[:, :, W - 1, :].astype(np.int64)
        t = (t << np.int64(64 - top_bits)) >> np.int64(64 - top_bits)
        w[:, :, W - 1, :] = t.astype(np.uint64)
    return w


def words_to_ints(words):
    """uint64[W][n] planar two's complement -> list of python ints."""
    W, n = words.shape
    out = []
    for i in range(n):
        v = 0
        for w in range(W):
            v |= int(words[w, i]) << (64 * w)
        if v >> (64 * W - 1):
            v -= 1 << (64 * W)
        out.append(v)
    return out


def ints_to_words(vals, W):
    n = len(vals)
    out = np.zeros((W, n), dtype=np.uint64)
    mask = (1 << 64) - 1
    for i, v in enumerate(vals):
        v &= (1 << (64 * W)) - 1
        for w in range(W):
            out[w, i] = (v >> (64 * w)) & mask
    return out


@pytest.fixture(scope="session")
def ref_available():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built (run oracle/build_ref.sh where /root/reference exists)")
    return True
