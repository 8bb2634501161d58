"""hb_he_mulpt_static: the allocation-free variant of hb_he_mulpt (scratch in the library's static device array).

No counterpart in the reference, so the check is against the existing
entry point: same inputs through hb_he_mulpt and hb_he_mulpt_static, called through the C ABI, must give the same
words bit for bit (integer arithmetic: the tolerance of every test here is zero) and the same nu / B -- on two calls
in a row that differ in batch size, level and data, so that the contents the first call leaves in the shared array
would show in the second.  The committed golden vector of the reference's he_mulpt pins the absolute values.

Collected last (file name), runs alone in a fresh process as well as after the other GPU tests: it builds its own
contexts and touches no module-level state of the other files."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, random_ct_words

pytestmark = pytest.mark.gpu
GOLD_DIR = os.path.join(os.path.dirname(__file__), "golden")
A_PARAMS = (12, 109, 16, 50)      # config A: cstr-hempc (src/ctr.c:510-518)


def _pair(eng, a_words, l, nu, pt):
    """(existing entry point, static entry point) on the same uploaded ciphertext."""
    a = eng.ct(a_words, l, nu=nu, B=3.0)
    out = eng.ct(batch=a.batch)
    n0 = eng.launches()
    want = eng.he_mulpt(a, pt)
    n1 = eng.launches()
    got = eng.he_mulpt_static(a, pt, out)
    return want, got, (n1 - n0, eng.launches() - n1)


def test_static_equals_he_mulpt_two_calls_in_a_row():
    from hectr_b200 import Engine
    eng = Engine(*A_PARAMS)
    try:
        rng = np.random.default_rng(2025)
        launches = []
        # first call: a large batch at the top level; second: a small batch one level down with another plaintext
        for batch, l, nu in ((48, eng.L, 2.0 ** 50), (3, eng.L - 1, 2.0 ** 50), (5, 1, 2.0 ** 50)):
            m = rng.standard_normal(16) + 1j * rng.standard_normal(16)
            pt = eng.pt(eng.encode(m), 2.0 ** 50)
            a = random_ct_words(rng, batch, eng.words(l), eng.n, eng.logq(l))
            want, got, n = _pair(eng, a, l, nu, pt)
            launches.append(n)
            w, g = want.download(), got.download()
            assert w.shape == g.shape == (batch, 2, eng.words(l), eng.n)
            assert np.array_equal(w, g), (batch, l)
            assert (got.l, got.nu, got.B) == (want.l, want.nu, want.B)
            assert np.any(g != 0)
            eng.pt_free(pt)
        assert all(n[0] == n[1] and n[1] > 0 for n in launches), launches       # the same kernels both ways
        # the same call again, in place (dst aliases src), after the array has been used by other shapes
        m = rng.standard_normal(16)
        pt = eng.pt(eng.encode(m + 0j), 2.0 ** 50)
        a = random_ct_words(rng, 7, eng.words(eng.L), eng.n, eng.logq(eng.L))
        want = eng.he_mulpt(eng.ct(a, eng.L), pt).download()
        c = eng.ct(a, eng.L)
        for _ in range(2):
            c = eng.ct(a, eng.L)
            assert np.array_equal(eng.he_mulpt_static(c, pt, c).download(), want)
    finally:
        eng.close()


def test_static_reproduces_the_reference_golden_vector():
    from hectr_b200 import Engine
    gold = np.load(os.path.join(GOLD_DIR, "small_logn7.npz"))
    eng = Engine(*[int(v) for v in gold["params"]])
    try:
        a = eng.ct(gold["ct_in"][0], eng.L)
        pt = eng.pt(gold["pt_words"], 2.0 ** 50)
        for _ in range(2):
            out = eng.ct(batch=1)
            assert np.array_equal(eng.he_mulpt_static(a, pt, out).download()[0], gold["mulpt"])
    finally:
        eng.close()


_OVERSIZE = r"""
import numpy as np
from hectr_b200 import Engine
eng = Engine(12, 109, 16, 50)
pt = eng.pt(eng.encode(np.ones(16, dtype=np.complex128)), 2.0 ** 50)
a = eng.ct(batch=1024)
a.upload(np.zeros((1024, 2, eng.words(eng.L), eng.n), dtype=np.uint64), eng.L, 1.0, 1.0)
out = eng.ct(batch=1024)
print("calling", flush=True)
eng.he_mulpt_static(a, pt, out)
print("returned", flush=True)
"""


def test_static_rejects_a_batch_beyond_the_array():
    """Like every misuse of the C ABI: a red `error:` line and abort()."""
    r = subprocess.run([sys.executable, "-c", _OVERSIZE], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert "calling" in r.stdout and "returned" not in r.stdout, (r.stdout, r.stderr[-400:])
    assert r.returncode != 0
    assert "hb_he_mulpt_static" in r.stderr and "bytes of scratch" in r.stderr, r.stderr[-400:]
