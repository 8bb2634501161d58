"""Pins the oracle (CPU, no GPU): the big-integer restatement ``oracle/restate.py`` against
  (1) the reference's own known answers -- GPQHE/tests/polymul.gp:1-23 (the two logn=7 primes appear literally in
      that script; its product is the exact negacyclic product mod 2^61, centred, recomputed here by schoolbook), and
  (2) tests/golden/small_logn7.npz, vectors written by the compiled UNMODIFIED reference (tests/golden/make_golden.py).
and, when oracle/_ref is present, the compiled reference against the same fixtures (they must not drift)."""
import os

import numpy as np
import pytest

from conftest import ints_to_words, words_to_ints
from oracle import restate as R

GOLD = os.path.join(os.path.dirname(__file__), "golden", "small_logn7.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def ctx(gold):
    return R.Ctx(*[int(v) for v in gold["params"]])


def ct_of(words, l):
    return {"l": l, "c0": words_to_ints(words[0]), "c1": words_to_ints(words[1])}


def words_of(ct, W):
    return np.stack([ints_to_words(ct["c0"], W), ints_to_words(ct["c1"], W)])


def evk_of(arr):
    return ([list(map(int, arr[0, d])) for d in range(arr.shape[1])], [list(map(int, arr[1, d])) for d in range(arr.shape[1])])


def test_polymul_gp_primes_and_product():
    """GPQHE/tests/polymul.gp: q = 2^61, ring X^128+1; a_i = i+2, b_i = i+3, and the second case whose
    coefficients are built from the two logn=7 chain primes 576460752303434497, 576460752303436801."""
    c = R.Ctx(7, 61, 16, 50)
    assert [p.p for p in c.primes[:2]] == [576460752303434497, 576460752303436801]
    q = 1 << 61
    for a, b in (([i + 2 for i in range(128)], [i + 3 for i in range(128)]),
                 ([576460752303434497 - i - 1 for i in range(128)], [576460752303436801 - i - 1 for i in range(128)])):
        want = [R.smod(v, q) for v in R.negacyclic_schoolbook(a, b)]
        # poly_mul with enough limbs for the exact product (|coeff| < 128 * 2^60 * 2^60 < P_3 / 2)
        assert c.poly_mul(a, b, 3, q) == want


def test_scalar_semantics():
    q = 1 << 10
    assert R.smod(511, q) == 511 and R.smod(512, q) == -512 and R.smod(-513, q) == 511      # types.c:108-113
    assert R.rdiv(5, 2) == 2 and R.rdiv(7, 2) == 3 and R.rdiv(-5, 2) == -3 and R.rdiv(6, 4) == 1 and R.rdiv(7, 4) == 2
    assert R.rdiv(-6, 4) == -2 and R.rdiv(-7, 4) == -2                                          # floor + [r > m/2]


def test_tables_match_golden(ctx, gold):
    assert [p.p for p in ctx.primes] == [int(v) for v in gold["primes"]]
    assert [ctx.L, ctx.dim, ctx.dimevk, ctx.dimub] == [int(v) for v in gold["derived"]]
    assert ctx.primes[0].zetas == [int(v) for v in gold["zetas0"]]
    assert ctx.primes[0].zetas_inv == [int(v) for v in gold["zetas_inv0"]]


def test_transforms_match_golden(ctx, gold):
    for x, y in zip(gold["ntt_in"], gold["ntt_out"]):
        assert ctx.ntt([int(v) for v in x], 0) == [int(v) for v in y]
    for x, y in zip(gold["intt_in"], gold["intt_out"]):
        assert ctx.invntt([int(v) for v in x], 0) == [int(v) for v in y]
    p0 = ctx.primes[0].p
    assert int(gold["ntt_out"][3].max()) <= p0           # the [0,p] representation is reproduced above


def test_elementwise_ops_match_golden(ctx, gold):
    l = ctx.L
    a, b = ct_of(gold["ct_in"][0], l), ct_of(gold["ct_in"][1], l)
    assert np.array_equal(words_of(ctx.he_add(a, b), 1), gold["add"])
    assert np.array_equal(words_of(ctx.he_sub(a, b), 1), gold["sub"])
    assert np.array_equal(words_of(ctx.he_neg(a), 1), gold["neg"])
    assert np.array_equal(words_of(ctx.he_rs(a), gold["rs"].shape[1]), gold["rs"])
    assert np.array_equal(words_of(ctx.he_moddown(a), gold["moddown"].shape[1]), gold["moddown"])


def test_mulpt_rot_conj_mul_match_golden(ctx, gold):
    l = ctx.L
    a, b = ct_of(gold["ct_in"][0], l), ct_of(gold["ct_in"][1], l)
    pt = words_to_ints(gold["pt_words"])
    assert np.array_equal(words_of(ctx.he_mulpt(a, pt, float(ctx.logDelta)), 1), gold["mulpt"])
    rk = [evk_of(k) for k in gold["rk"]]
    for r in (0, 1, 5, 15):
        assert np.array_equal(words_of(ctx.he_rot(a, r, rk), 1), gold["rot"][r]), r
    assert np.array_equal(words_of(ctx.he_conj(a, evk_of(gold["ck"])), 1), gold["conj"])
    assert np.array_equal(words_of(ctx.he_mul(a, b, evk_of(gold["rlk"])), 1), gold["mul"])


def test_gemv_and_hempc_match_golden(ctx, gold):
    l = ctx.L
    rk = [evk_of(k) for k in gold["rk"]]
    cts = [ct_of(w, l) for w in gold["ct_in"]]
    Ka_pts = [words_to_ints(w) for w in gold["Ka_pts"]]
    Kb_pts = [words_to_ints(w) for w in gold["Kb_pts"]]
    assert np.array_equal(words_of(ctx.he_gemv(cts[0], Ka_pts, rk), gold["gemv"].shape[1]), gold["gemv"])
    up = ctx.ctr_hempc(*cts, Ka_pts, Kb_pts, rk)
    assert np.array_equal(words_of(up, gold["hempc"].shape[1]), gold["hempc"])


def test_compiled_reference_reproduces_golden(ref_available, gold):
    """The fixture generator is deterministic: the compiled reference still produces the committed vectors."""
    from oracle.ref import Ref
    ref = Ref.get(*[int(v) for v in gold["params"]])
    assert ref.primes() == [int(v) for v in gold["primes"]]
    for x, y in zip(gold["ntt_in"], gold["ntt_out"]):
        assert np.array_equal(ref.ntt(x, 0), y)
    a = ref.ct_from_words(gold["ct_in"][0], ref.L)
    b = ref.ct_from_words(gold["ct_in"][1], ref.L)
    assert np.array_equal(ref.ct_to_words(ref.add(a, b)), gold["add"])
    assert np.array_equal(ref.ct_to_words(ref.rs(ref.copy_ct(a))), gold["rs"])
