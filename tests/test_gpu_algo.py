"""GPU parity for he_addpt / he_subpt / he_const_pt and the he-algo.c compositions against the
compiled, unmodified reference.  A deep level chain (q = 2^520, Delta = 2^40, L = 13) on the reference's small KAT ring
(N = 128: below 2^10 the reference accepts any q, precomp.c:339-340) keeps the CPU side to seconds; every function is a
fixed sequence of he_mul / he_rs / he_moddown / he_mulpt / he_addpt ..., so a bit-exact result pins the order of
operations, the level bookkeeping and every primitive across thirteen levels (1 to 9 words per coefficient)."""
import ctypes as C

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


@pytest.mark.parametrize("alpha", [1, 2])
def test_cmppt(deep, alpha):
    """alpha = 2 would mean two squaring rounds in he_cmp; he_cmppt's unsigned negation (he-algo.c:537) makes it none"""
    ref, eng = deep["ref"], deep["eng"]
    want = ref.he_cmppt(deep["ct1"], ref.encode(deep["m2"]), deep["rlk"], 1, alpha)
    assert want.l == ref.L - 4
    pt = eng.he_ecd(deep["m2"])
    got = eng.he_cmppt(_up(deep, deep["ct1"]), pt, deep["k_rlk"], 1, alpha)
    _same(deep, got, want)
    eng.pt_free(pt)


# ---- key generation on the device ------------------------------------------------------------------
def _genswk_case(ref, eng, sk, sps):
    """sps: int64[nkeys][N] polynomials switched from.  Draws e and a from the reference's samplers in he_genswk's order,
    runs the device, and rebuilds the expected limbs from the reference's own poly_mul / rns_decompose / ntt."""
    from conftest import ints_to_words, words_to_ints
    n, nkeys = ref.n, len(sps)
    P, M = ref.mpi_int(ref.hectx.P), ref.mpi_int(ref.hectx.PqL)
    dimK = (M.bit_length() + ref.logn) // 59 + 1
    WA = eng.genswk_words()
    assert WA == (M.bit_length() + 1 + 63) // 64
    es, a_polys = [], []
    for _ in range(nkeys):
        es.append(ref.sample_error())                       # he-kem.c:89
        a_polys.append(ref.sample_uniform_PqL())             # he-kem.c:94
    a_words = np.stack([ref.poly_to_words(a, WA) for a in a_polys])
    dsk = eng.sk(ref.poly_to_words(sk, 1)[0].view(np.int64))
    got = eng.he_genswk(a_words, np.stack(es), np.stack(sps), dsk)
    smod = lambda x: ((x % M) - M) if (x % M) >= M // 2 else (x % M)
    tmp = ref.alloc_poly()
    for k in range(nkeys):
        x = ref.poly_mul_PqL(a_polys[k], sk, dimK)
        a_int = words_to_ints(a_words[k])
        p0 = [smod(-x[i] + int(es[k][i]) + P * int(sps[k][i])) for i in range(n)]
        p1 = [smod(v) for v in a_int]
        for h, vals in enumerate((p0, p1)):
            ref.words_to_poly(tmp, ints_to_words(vals, WA))
            for d in range(ref.dimevk):
                want = ref.ntt(ref.rns_decompose(tmp, d), d)
                assert np.array_equal(got[k, h, d], want), (k, h, d)
    return got, dsk


def test_genswk_deep_chain(deep):
    """rotation-key and conjugation-key shaped inputs on the 13-level ring (P q_L of 9+ words)"""
    ref, eng, sk = deep["ref"], deep["eng"], deep["sk"]
    s = ref.poly_to_words(sk, 1)
    sps = [ref.poly_rot(s, r)[0].view(np.int64) for r in (0, 3)]
    _genswk_case(ref, eng, sk, sps)


def test_genswk_config_A_rotation_keys_work(ref_available):
    """Config A: generate rk[1], rk[5] and the relinearisation key's shape (sp = sk^2 mod q_L) on the device, bit-exact vs
    the reference's arithmetic; then rotate a reference ciphertext with the device-made key and decrypt it with the reference."""
    from oracle.ref import Ref
    from hectr_b200 import Engine
    params = (12, 109, 16, 50)
    ref = Ref.get(*params)
    eng = Engine(*params)
    try:
        pk, sk = ref.keypair()
        s = ref.poly_to_words(sk, 1)
        L = ref.L
        W = eng.words(L)
        s2 = ref.poly_mul_words(sk, s, eng.logq(L) // 59 + 1, eng.logq(L), W)            # he_genrlk, he-kem.c:133
        # |sk^2| <= hamming weight: one signed word; the higher words of the W-word form are its sign extension
        assert all(np.array_equal(s2[w].view(np.int64), s2[0].view(np.int64) >> 63) for w in range(1, W))
        sps = [ref.poly_rot(s, 1)[0].view(np.int64), ref.poly_rot(s, 5)[0].view(np.int64), s2[0].view(np.int64)]
        got, dsk = _genswk_case(ref, eng, sk, sps)
        keys = np.zeros((16,) + got.shape[1:], dtype=np.uint64)
        keys[1], keys[5] = got[0], got[1]
        rk = eng.keys(keys)
        m = np.arange(1, 17) * (0.25 + 0.125j)
        ct = ref.enc_pk(ref.encode(m), pk)
        for r in (1, 5):
            out = eng.he_rot(eng.ct(ref.ct_to_words(ct), L, nu=ct.nu, B=ct.B), r, rk)
            dec = ref.decode(ref.dec(ref.ct_from_words(out.download()[0], L, nu=out.nu, B=out.B), sk))
            assert np.abs(dec - np.roll(m, -r)).max() < 1e-6, r
        rk.free()
    finally:
        eng.close()


GENRK_CHILD = r"""
import sys
sys.path.insert(0, %(root)r)
import numpy as np
from oracle.ref import Ref
mode, out = sys.argv[1], sys.argv[2]
params = (12, 109, 16, 50)
ref = Ref.get(*params)
pk, sk = ref.keypair()
NK = 3
if mode == "ref":
    rk = ref.genrk(sk)                                   # he_genrk itself (he-kem.c:154-169)
    np.save(out, np.stack([ref.evk_to_array(rk[i]) for i in range(NK)]))
else:
    from hectr_b200 import Engine
    eng = Engine(*params)
    s = ref.poly_to_words(sk, 1)
    WA = eng.genswk_words()
    es, aws, sps = [], [], []
    for k in range(NK):                                  # he_genswk's draws, in its order (he-kem.c:89,94)
        sps.append(ref.poly_rot(s, k)[0].view(np.int64))
        es.append(ref.sample_error())
        aws.append(ref.poly_to_words(ref.sample_uniform_PqL(), WA))
    dsk = eng.sk(s[0].view(np.int64))
    np.save(out, eng.he_genswk(np.stack(aws), np.stack(es), np.stack(sps), dsk))
    eng.close()
"""


def test_genswk_equals_he_genrk_itself(ref_available, tmp_path):
    """hb_he_genswk against the keys he_genrk ITSELF emits for the same SUPERCOP stream: he_genswk is static and samples
    internally, but the stream restarts with every process, so one fresh process runs he_keypair + he_genrk and another
    he_keypair + the same draws by hand + the device."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for mode in ("ref", "dev"):
        out = str(tmp_path / (mode + ".npy"))
        r = subprocess.run([sys.executable, "-c", GENRK_CHILD % {"root": root}, mode, out], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(out))
    assert outs[0].shape == outs[1].shape
    assert np.array_equal(outs[0], outs[1])


@pytest.mark.parametrize("rot", [0, 1, 7, -1])
def test_poly_rot_direct(deep, rot):
    """hb_poly_rot against poly_rot / poly_conj (GPQHE/src/poly.c:263-283): rotation 0, 1, slots-1 and conjugation (rot < 0),
    multiword coefficients including the negation edge -2^(k-1) and zero."""
    from conftest import random_ct_words
    from oracle.ref import Ref
    eng = deep["eng"]
    ref = Ref.get(*PARAMS)          # the reference keeps ONE global context: an earlier test of this module re-initialised it
    rng = np.random.default_rng(300 + rot)
    L = ref.L
    W, k = eng.words(L), eng.logq(L)
    w = random_ct_words(rng, 1, W, eng.n, k)[0]           # [2][W][N]
    w[0, :, 0] = 0
    w[0, :, 1] = 0
    w[0, W - 1, 1] = (1 << 64) - (1 << ((k - 1) % 64))     # -2^(k-1)
    got = eng.poly_rot(w, rot)
    for q in range(2):
        if rot >= 0:
            want = ref.poly_rot(w[q], rot)
        else:
            a, r = ref.alloc_poly(), ref.alloc_poly()
            ref.words_to_poly(a, w[q])
            ref.lib.poly_conj(C.byref(r), C.byref(a))
            want = ref.poly_to_words(r, W)
        assert np.array_equal(got[q], want), (rot, q)
