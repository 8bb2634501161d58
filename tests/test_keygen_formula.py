"""Pins the expected-value formula the GPU key-generation test uses (tests/test_gpu_algo.py::_genswk_case) against the
reference's own he_genrk.  he_genswk is static and draws its samples internally; the reference's SUPERCOP stream
(GPQHE/src/rng.c:32-78) is deterministic per process, so two fresh processes that issue the same calls see the same
draws: one runs he_keypair + he_genrk, the other he_keypair and then the draws of he_genswk by hand
(sample_error, sample_uniform: he-kem.c:89,94) followed by the formula built from poly_mul / rns_decompose / ntt."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys
sys.path.insert(0, %(root)r)
sys.path.insert(0, %(root)r + "/tests")
import numpy as np
from conftest import ints_to_words, words_to_ints
from oracle.ref import Ref
mode, out = sys.argv[1], sys.argv[2]
ref = Ref.get(7, 150, 8, 40)
pk, sk = ref.keypair()
NK = 3
if mode == "ref":
    rk = ref.genrk(sk)
    np.save(out, np.stack([ref.evk_to_array(rk[i]) for i in range(NK)]))
else:
    n = ref.n
    P, M = ref.mpi_int(ref.hectx.P), ref.mpi_int(ref.hectx.PqL)
    dimK = (M.bit_length() + ref.logn) // 59 + 1
    WA = (M.bit_length() + 1 + 63) // 64
    s = ref.poly_to_words(sk, 1)
    smod = lambda x: ((x %% M) - M) if (x %% M) >= M // 2 else (x %% M)
    res = np.zeros((NK, 2, ref.dimevk, n), dtype=np.uint64)
    tmp = ref.alloc_poly()
    for k in range(NK):
        sp = ref.poly_rot(s, k)[0].view(np.int64)
        e = ref.sample_error()
        a = ref.sample_uniform_PqL()
        x = ref.poly_mul_PqL(a, sk, dimK)
        a_int = words_to_ints(ref.poly_to_words(a, WA))
        p0 = [smod(-x[i] + int(e[i]) + P * int(sp[i])) for i in range(n)]
        p1 = [smod(v) for v in a_int]
        for h, vals in enumerate((p0, p1)):
            ref.words_to_poly(tmp, ints_to_words(vals, WA))
            for d in range(ref.dimevk):
                res[k, h, d] = ref.ntt(ref.rns_decompose(tmp, d), d)
    np.save(out, res)
"""


def test_genswk_formula_matches_he_genrk(ref_available, tmp_path):
    outs = []
    for mode in ("ref", "formula"):
        out = str(tmp_path / (mode + ".npy"))
        r = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}, mode, out], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(out))
    assert outs[0].shape == outs[1].shape
    assert np.array_equal(outs[0], outs[1])
