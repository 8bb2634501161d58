"""GPU parity against the COMMITTED golden vectors (tests/golden/, written by the compiled unmodified reference through
tests/golden/make_golden.py).  Needs no oracle at run time: the CUDA path, called through the C ABI, must reproduce
the reference's words bit for bit."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import random_ct_words

pytestmark = pytest.mark.gpu
GOLD_DIR = os.path.join(os.path.dirname(__file__), "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD_DIR, "small_logn7.npz"))


@pytest.fixture(scope="module")
def small(gold):
    from hectr_b200 import Engine
    eng = Engine(*[int(v) for v in gold["params"]])
    rk = eng.keys(gold["rk"])
    rlk = eng.keys(gold["rlk"])
    ck = eng.keys(gold["ck"])
    yield dict(eng=eng, rk=rk, rlk=rlk, ck=ck)
    for k in (rk, rlk, ck):
        k.free()
    eng.close()


def test_small_tables(small, gold):
    eng = small["eng"]
    assert list(map(int, eng.primes())) == [int(v) for v in gold["primes"]]
    i = eng.info
    assert [i.L, i.dim, i.dimevk, i.dimub] == [int(v) for v in gold["derived"]]
    # the engine keeps plain powers zeta^brv(i); the reference stores them in Montgomery form, * 2^64 mod p (precomp.c:259-260)
    p0 = int(gold["primes"][0])
    assert [int(v) * (1 << 64) % p0 for v in eng.twiddles(0)] == [int(v) for v in gold["zetas0"]]
    assert [int(v) * (1 << 64) % p0 for v in eng.twiddles(0, inverse=True)] == [int(v) for v in gold["zetas_inv0"]]


def test_small_transforms(small, gold):
    eng = small["eng"]
    assert np.array_equal(eng.ntt(gold["ntt_in"], strict=True), gold["ntt_out"])       # [0,p] representation, SURVEY F8
    p0 = np.uint64(int(gold["primes"][0]))
    assert np.array_equal(eng.ntt(gold["ntt_in"]), gold["ntt_out"] % p0)                # default: canonical
    assert np.array_equal(eng.ntt(gold["intt_in"], inverse=True), gold["intt_out"])


def test_small_elementwise(small, gold):
    eng = small["eng"]
    l = eng.L
    a, b = eng.ct(gold["ct_in"][0], l), eng.ct(gold["ct_in"][1], l)
    assert np.array_equal(eng.he_add(a, b).download()[0], gold["add"])
    assert np.array_equal(eng.he_sub(a, b).download()[0], gold["sub"])
    assert np.array_equal(eng.he_neg(eng.he_copy_ct(a)).download()[0], gold["neg"])
    assert np.array_equal(eng.he_rs(eng.he_copy_ct(a)).download()[0], gold["rs"])
    assert np.array_equal(eng.he_moddown(eng.he_copy_ct(a)).download()[0], gold["moddown"])


def test_small_mulpt_rot_conj_mul(small, gold):
    eng = small["eng"]
    l = eng.L
    a, b = eng.ct(gold["ct_in"][0], l), eng.ct(gold["ct_in"][1], l)
    pt = eng.pt(gold["pt_words"], 2.0 ** 50)
    assert np.array_equal(eng.he_mulpt(a, pt).download()[0], gold["mulpt"])
    for r in range(eng.slots):
        assert np.array_equal(eng.he_rot(eng.he_copy_ct(a), r, small["rk"]).download()[0], gold["rot"][r]), r
    assert np.array_equal(eng.he_conj(eng.he_copy_ct(a), small["ck"]).download()[0], gold["conj"])
    assert np.array_equal(eng.he_mul(a, b, small["rlk"]).download()[0], gold["mul"])


def test_small_gemv_and_hempc(small, gold):
    eng = small["eng"]
    l = eng.L
    cts = [eng.ct(w, l) for w in gold["ct_in"]]
    assert np.array_equal(eng.he_gemv(gold["Ka"], cts[0], small["rk"]).download()[0], gold["gemv"])
    up = eng.ctr_hempc_eval(gold["Ka"], gold["Kb"], *cts, small["rk"])
    assert np.array_equal(up.download()[0], gold["hempc"])
    # a batch of identical instances gives identical instances (the batched launch is instance-independent)
    stacked = [eng.ct(np.stack([w, w, w]), l) for w in gold["ct_in"]]
    upb = eng.ctr_hempc_eval(gold["Ka"], gold["Kb"], *stacked, small["rk"]).download()
    for i in range(3):
        assert np.array_equal(upb[i], gold["hempc"])
    # decoded control action: the host decoder on the decrypted plaintext is not on the device path; compare the
    # ciphertext-level result only, and check the fixture carries the decoded reference value for documentation
    assert gold["hempc_decoded"].shape == (eng.slots,)


def _synth_keys(rng, primes, nkeys, dimevk, n):
    keys = np.empty((nkeys, 2, dimevk, n), dtype=np.uint64)
    for d in range(dimevk):
        keys[:, :, d, :] = rng.integers(0, int(primes[d]), size=(nkeys, 2, n), dtype=np.uint64)
    return keys


@pytest.mark.parametrize("tag", ["A", "B13"])
def test_digests_of_reference_results(tag):
    """Inputs are a pure function of a numpy seed; tests/golden/digests.json holds SHA-256 of the reference's outputs."""
    from hectr_b200 import Engine
    d = json.load(open(os.path.join(GOLD_DIR, "digests.json")))[tag]
    eng = Engine(*d["params"])
    rng = np.random.default_rng(d["seed"])
    n, S, l = eng.n, eng.slots, eng.L
    W, k = eng.words(l), eng.logq(l)
    keys = _synth_keys(rng, eng.primes(), S + 1, eng.info.dimevk, n)
    w = random_ct_words(rng, 4, W, n, k)
    assert sha(w) == d["inputs"] and sha(keys) == d["keys"], "numpy stream changed: regenerate tests/golden"
    rk, rlk = eng.keys(keys[:S]), eng.keys(keys[S])
    cts = [eng.ct(w[i], l) for i in range(4)]
    assert sha(eng.he_sub(cts[0], cts[1]).download()[0]) == d["sub"]
    assert sha(eng.he_rot(eng.he_copy_ct(cts[0]), 3, rk).download()[0]) == d["rot3"]
    assert sha(eng.he_rs(eng.he_copy_ct(cts[0])).download()[0]) == d["rs"]
    assert sha(eng.he_mul(cts[0], cts[1], rlk).download()[0]) == d["mul"]
    if "gemv" in d:
        Ka = np.zeros((S, S), dtype=np.complex128)
        Kb = np.zeros((S, S), dtype=np.complex128)
        Ka[:8, :3] = rng.standard_normal((8, 3))
        Kb[:8, :2] = rng.standard_normal((8, 2))
        assert sha(eng.he_gemv(Ka, cts[0], rk).download()[0]) == d["gemv"]
        assert sha(eng.ctr_hempc_eval(Ka, Kb, *cts, rk).download()[0]) == d["hempc"]
    rk.free()
    rlk.free()
    eng.close()
