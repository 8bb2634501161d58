"""GPU parity for he_addpt / he_subpt / he_const_pt and the he-algo.c compositions against the
compiled, unmodified reference.  A deep level chain (q = 2^520, Delta = 2^40, L = 13) on the reference's small KAT ring
(N = 128: below 2^10 the reference accepts any q, precomp.c:339-340) keeps the CPU side to seconds; every function is a
fixed sequence of he_mul / he_rs / he_moddown / he_mulpt / he_addpt ..., so a bit-exact result pins the order of
operations, the level bookkeeping and every primitive across thirteen levels (1 to 9 words per coefficient)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
PARAMS = (7, 520, 8, 40)


@pytest.fixture(scope="module")
def deep(ref_available):
    from oracle.ref import Ref
    from hectr_b200 import Engine
    ref = Ref.get(*PARAMS)
    eng = Engine(*PARAMS)
    pk, sk = ref.keypair()
    rlk, ck, rk = ref.genrlk(sk), ref.genck(sk), ref.genrk(sk)
    k_rlk = eng.keys(ref.evk_to_array(rlk))
    k_ck = eng.keys(ref.evk_to_array(ck))
    k_rk = eng.keys(np.stack([ref.evk_to_array(rk[i]) for i in range(ref.slots)]))
    m1 = np.array([0.5, 0.8, 1.2, 0.9, 1.1, 0.7, 0.6, 1.3], dtype=np.complex128)
    m2 = np.array([0.9, 0.4, 1.0, 1.4, 0.3, 0.7, 1.1, 0.2], dtype=np.complex128)
    ct1, ct2 = ref.enc_pk(ref.encode(m1), pk), ref.enc_pk(ref.encode(m2), pk)
    yield dict(ref=ref, eng=eng, sk=sk, rlk=rlk, ck=ck, rk=rk, k_rlk=k_rlk, k_ck=k_ck, k_rk=k_rk, m1=m1, m2=m2, ct1=ct1, ct2=ct2)
    for k in (k_rlk, k_ck, k_rk):
        k.free()
    eng.close()


def _up(d, ct):
    """reference ciphertext -> engine ciphertext (batch of 2 identical instances: the batched launch must not mix them)"""
    w = d["ref"].ct_to_words(ct)
    return d["eng"].ct(np.stack([w, w]), ct.l, nu=ct.nu, B=ct.B)


def _same(d, got, want):
    ref = d["ref"]
    assert got.l == want.l
    out = got.download()
    w = ref.ct_to_words(want)
    assert np.array_equal(out[0], w) and np.array_equal(out[1], w)
    assert got.nu == want.nu


def test_const_pt_addpt_subpt(deep):
    ref, eng = deep["ref"], deep["eng"]
    for num in (1.0, 0.25, -1. / 48, 3 + 2j, (1. / 480) / (-17. / 80640)):
        want = ref.const_pt(num)
        got = eng.he_const_pt(num)
        W = eng.info.words_top
        host = eng.from_device(got.words, (W, eng.n))
        assert np.array_equal(host, ref.poly_to_words(want.m, W)), num
        for l in (ref.L, 3):
            a = deep["ct1"]
            a_l = ref.copy_ct(a)
            for _ in range(ref.L - l):
                ref.moddown(a_l)
            ga = _up(deep, a_l)
            _same(deep, eng.he_addpt(ga, got), ref.addpt(a_l, want))
            _same(deep, eng.he_subpt(ga, got), ref.subpt(a_l, want))
        eng.pt_free(got)


def test_sum_idx_nrm2(deep):
    ref, eng = deep["ref"], deep["eng"]
    a = _up(deep, deep["ct1"])
    _same(deep, eng.he_sum(a, deep["k_rk"]), ref.he_sum(deep["ct1"], deep["rk"]))
    _same(deep, eng.he_idx(a, 3, deep["k_rk"]), ref.he_idx(deep["ct1"], 3, deep["rk"]))
    _same(deep, eng.he_nrm2(a, deep["k_rlk"], deep["k_ck"], deep["k_rk"]), ref.he_nrm2(deep["ct1"], deep["rlk"], deep["ck"], deep["rk"]))


@pytest.mark.parametrize("iters", [1, 4])
def test_inv(deep, iters):
    ref, eng = deep["ref"], deep["eng"]
    want = ref.he_inv(deep["ct1"], deep["rlk"], iters)
    got = eng.he_inv(_up(deep, deep["ct1"]), deep["k_rlk"], iters)
    _same(deep, got, want)
    if iters == 4:
        dec = ref.decode(ref.dec(want, deep["sk"]))
        assert np.allclose(dec, 1 / deep["m1"], atol=2e-2)          # the function it computes, for the record


def test_sqrt(deep):
    ref, eng = deep["ref"], deep["eng"]
    _same(deep, eng.he_sqrt(_up(deep, deep["ct1"]), deep["k_rlk"], 3), ref.he_sqrt(deep["ct1"], deep["rlk"], 3))


def test_sigmoid_log(deep):
    ref, eng = deep["ref"], deep["eng"]
    a = _up(deep, deep["ct1"])
    _same(deep, eng.he_sigmoid(a, deep["k_rlk"]), ref.he_sigmoid(deep["ct1"], deep["rlk"]))
    _same(deep, eng.he_log(a, deep["k_rlk"]), ref.he_log(deep["ct1"], deep["rlk"]))


@pytest.mark.parametrize("coef,iters", [(1.0, 2), (-0.5 + 0.25j, 1)])
def test_exp(deep, coef, iters):
    ref, eng = deep["ref"], deep["eng"]
    _same(deep, eng.he_exp(coef, _up(deep, deep["ct1"]), deep["k_rlk"], iters), ref.he_exp(coef, deep["ct1"], deep["rlk"], iters))


@pytest.mark.parametrize("iters,alpha", [(1, 1), (1, 2)])
def test_cmp(deep, iters, alpha):
    """depth (3 + iter)(1 + t): 4 levels for alpha = 1 (t = 0), 12 for alpha = 2 (t = 2) of the 13 available"""
    ref, eng = deep["ref"], deep["eng"]
    want = ref.he_cmp(deep["ct1"], deep["ct2"], deep["rlk"], iters, alpha)
    got = eng.he_cmp(_up(deep, deep["ct1"]), _up(deep, deep["ct2"]), deep["k_rlk"], iters, alpha)
    _same(deep, got, want)
