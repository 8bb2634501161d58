"""GPU parity: the CUDA path (through the C ABI) against the compiled, unmodified reference
(oracle/_ref/libgpqhe.so) on the same seeded inputs.  Integer work: bit-exact."""
import ctypes as C
import json

import numpy as np
import pytest

from conftest import random_ct_words

pytestmark = pytest.mark.gpu

A_PARAMS = (12, 109, 16, 50)      # config A: cstr-hempc (src/ctr.c:510-518)


@pytest.fixture(scope="module")
def cfgA(ref_available):
    from oracle.ref import Ref
    from hectr_b200 import Engine
    ref = Ref.get(*A_PARAMS)
    eng = Engine(*A_PARAMS)
    pk, sk = ref.keypair()
    rk = ref.genrk(sk)
    rk_arr = np.stack([ref.evk_to_array(rk[i]) for i in range(ref.slots)])
    keys = eng.keys(rk_arr)
    yield dict(ref=ref, eng=eng, pk=pk, sk=sk, rk=rk, keys=keys)
    keys.free()
    eng.close()


def _ref_ct(ref, words, l):
    return ref.ct_from_words(words, l)


def test_context_matches_reference(cfgA):
    ref, eng = cfgA["ref"], cfgA["eng"]
    i = eng.info
    assert (i.L, i.dim, i.dimevk, i.dimub) == (ref.L, ref.dim, ref.dimevk, ref.dimub)
    assert list(map(int, eng.primes())) == ref.primes()
    for l in range(ref.L + 1):
        assert 1 << eng.logq(l) == ref.q_level(l)


def test_decompose_ntt(cfgA):
    ref, eng = cfgA["ref"], cfgA["eng"]
    rng = np.random.default_rng(7)
    for l, D in ((2, 7), (1, 3)):
        W, k = eng.words(l), eng.logq(l)
        words = random_ct_words(rng, 1, W, eng.n, k)[0]          # [2][W][N]
        got = eng.poly_ntt(words, D)
        for q in range(2):
            poly = ref.alloc_poly()
            ref.words_to_poly(poly, words[q])
            for d in range(D):
                want = ref.ntt(ref.rns_decompose(poly, d), d) % np.uint64(ref.primes()[d])
                assert np.array_equal(got[q, d], want), (l, q, d)


@pytest.mark.parametrize("D,k,W", [(2, 59, 1), (3, 109, 2), (7, 109, 2), (7, 59, 1), (4, 109, 2)])
def test_reconstruct(cfgA, D, k, W):
    ref, eng = cfgA["ref"], cfgA["eng"]
    rng = np.random.default_rng(100 + D + k)
    primes = ref.primes()
    # valid products keep |x| < P_D/4; emulate by CRT-ing a bounded integer instead of random residues
    bits = sum(p.bit_length() for p in primes[:D]) - D - 3
    n = eng.n
    vals = [int.from_bytes(rng.bytes((bits + 7) // 8), "little") % (1 << bits) * (1 if i % 2 else -1) for i in range(n)]
    vals[0], vals[1], vals[2] = 0, 1, -1
    limbs = np.array([[v % p for v in vals] for p in primes[:D]], dtype=np.uint64)
    want = ref.rns2mpi(limbs, k, W)
    got = eng.rns2words(limbs[None], k, W)[0]
    assert np.array_equal(got, want)


def test_add_sub_neg_rs_moddown(cfgA):
    ref, eng = cfgA["ref"], cfgA["eng"]
    rng = np.random.default_rng(11)
    l = 2
    W, k = eng.words(l), eng.logq(l)
    a = random_ct_words(rng, 2, W, eng.n, k)
    b = random_ct_words(rng, 2, W, eng.n, k)
    # edge coefficients: extremes of the centred range and rounding ties for he_rs
    a[0, 0, :, 0] = [0, 1 << 44]; b[0, 0, :, 0] = [0, 1 << 44]          # +q/4 + q/4 wraps to -q/2
    a[0, 0, :, 1] = [0, (1 << 64) - (1 << 44)]                            # -q/2
    a[0, 0, :, 2] = [1 << 49, 0]                                          # exact tie: rounds down
    a[0, 0, :, 3] = [(1 << 49) + 1, 0]                                    # just above the tie
    a[0, 0, :, 4] = [(1 << 64) - (1 << 49), (1 << 64) - 1]                # -2^49: tie on a negative
    ea, eb = eng.ct(a, l), eng.ct(b, l)
    for bi in range(2):
        ra, rb = _ref_ct(ref, a[bi], l), _ref_ct(ref, b[bi], l)
        assert np.array_equal(eng.he_add(ea, eb).download()[bi], ref.ct_to_words(ref.add(ra, rb)))
        assert np.array_equal(eng.he_sub(ea, eb).download()[bi], ref.ct_to_words(ref.sub(ra, rb)))
        n1 = eng.he_neg(eng.he_copy_ct(ea)).download()[bi]
        assert np.array_equal(n1, ref.ct_to_words(ref.neg(ref.copy_ct(ra))))
        r1 = eng.he_rs(eng.he_copy_ct(ea))
        rr = ref.rs(ref.copy_ct(ra))
        assert r1.l == rr.l == 1 and np.array_equal(r1.download()[bi], ref.ct_to_words(rr))
        assert r1.nu == rr.nu and r1.B == rr.B
        m1 = eng.he_moddown(eng.he_copy_ct(ea))
        mr = ref.moddown(ref.copy_ct(ra))
        assert np.array_equal(m1.download()[bi], ref.ct_to_words(mr))


@pytest.mark.parametrize("l", [2, 1])
def test_rot_all_rotations(cfgA, l):
    ref, eng, keys = cfgA["ref"], cfgA["eng"], cfgA["keys"]
    rng = np.random.default_rng(21 + l)
    W, k = eng.words(l), eng.logq(l)
    a = random_ct_words(rng, 2, W, eng.n, k)
    a[0, :, :, 5] = [[0, (1 << 64) - (1 << 44)]] * 2 if W == 2 else [[(1 << 64) - (1 << 58)]] * 2   # -q/2 (negation edge)
    for r in range(ref.slots):
        got = eng.he_rot(eng.ct(a, l), r, keys).download()
        for bi in range(2):
            want = ref.ct_to_words(ref.rot(_ref_ct(ref, a[bi], l), r, cfgA["rk"]))
            assert np.array_equal(got[bi], want), (l, r, bi)


def test_rot_zero_c1_takes_exact_tail(cfgA):
    """c1 = 0 makes every key-switch product exactly 0 = 0 mod P: the fixed-point fast tail cannot decide R/P there and
    must fall back to the multiword path (k_ks_finish_fast -> ks_exact); the result is still the reference's."""
    ref, eng, keys = cfgA["ref"], cfgA["eng"], cfgA["keys"]
    rng = np.random.default_rng(77)
    l = 2
    a = random_ct_words(rng, 2, eng.words(l), eng.n, eng.logq(l))
    a[:, 1] = 0
    a[1, 0, :, ::3] = 0
    got = eng.he_rot(eng.ct(a, l), 3, keys).download()
    for bi in range(2):
        want = ref.ct_to_words(ref.rot(_ref_ct(ref, a[bi], l), 3, cfgA["rk"]))
        assert np.array_equal(got[bi], want)


def test_mulpt(cfgA):
    ref, eng = cfgA["ref"], cfgA["eng"]
    rng = np.random.default_rng(31)
    m = rng.standard_normal(16) + 1j * rng.standard_normal(16)
    pt_ref = ref.encode(m)
    pt_words = ref.poly_to_words(pt_ref.m, eng.info.words_top)
    assert np.array_equal(eng.encode(m), pt_words), "host encoder differs from he_ecd"
    pt = eng.pt(pt_words, pt_ref.nu)
    for l in (2, 1):
        a = random_ct_words(rng, 2, eng.words(l), eng.n, eng.logq(l))
        got = eng.he_mulpt(eng.ct(a, l, nu=2.0 ** 50), pt)
        for bi in range(2):
            ra = _ref_ct(ref, a[bi], l)
            ra.nu = 2.0 ** 50
            want = ref.mulpt(ra, pt_ref)
            assert np.array_equal(got.download()[bi], ref.ct_to_words(want)), (l, bi)
            assert got.nu == want.nu and got.B == want.B


def test_gemv_and_hempc(cfgA):
    ref, eng, keys, pk, sk = cfgA["ref"], cfgA["eng"], cfgA["keys"], cfgA["pk"], cfgA["sk"]
    rng = np.random.default_rng(41)
    S = 16
    Ka = np.zeros((S, S), dtype=np.complex128)
    Kb = np.zeros((S, S), dtype=np.complex128)
    Ka[:8, :3] = rng.standard_normal((8, 3))
    Kb[:8, :2] = rng.standard_normal((8, 2))
    cts = {}
    vals = {}
    for name in ("xhat", "xr", "uhat", "ur"):
        v = np.zeros(S, dtype=np.complex128)
        v[:3] = rng.standard_normal(3)
        vals[name] = v
        cts[name] = ref.enc_pk(ref.encode(v), pk)
    w = {n: ref.ct_to_words(c)[None] for n, c in cts.items()}
    e = {n: eng.ct(w[n], 2, nu=cts[n].nu, B=cts[n].B) for n in w}
    # gemv alone
    got = eng.he_gemv(Ka, e["xhat"], keys)
    want = ref.gemv(Ka, cts["xhat"], cfgA["rk"])
    assert got.l == want.l == 1
    assert np.array_equal(got.download()[0], ref.ct_to_words(want)), "he_gemv ciphertext differs"
    assert got.nu == want.nu and got.B == want.B
    # dense matrix (every diagonal non-zero, complex entries)
    Ad = rng.standard_normal((S, S)) + 1j * rng.standard_normal((S, S))
    gotd = eng.he_gemv(Ad, e["xhat"], keys)
    wantd = ref.gemv(Ad, cts["xhat"], cfgA["rk"])
    assert np.array_equal(gotd.download()[0], ref.ct_to_words(wantd)), "dense he_gemv ciphertext differs"
    # the ciphertext half of ctr_hempc (src/hempc.c:253-266)
    xd, ud = ref.sub(cts["xhat"], cts["xr"]), ref.sub(cts["uhat"], cts["ur"])
    du = ref.neg(ref.add(ref.gemv(Ka, xd, cfgA["rk"]), ref.gemv(Kb, ud, cfgA["rk"])))
    up_ref = ref.add(ref.moddown(ref.copy_ct(cts["uhat"])), du)
    up = eng.ctr_hempc_eval(Ka, Kb, e["xhat"], e["xr"], e["uhat"], e["ur"], keys)
    assert np.array_equal(up.download()[0], ref.ct_to_words(up_ref)), "ctr_hempc ciphertext differs"
    out = eng.ctr_hempc_eval_host(Ka, Kb, w["xhat"], w["xr"], w["uhat"], w["ur"], keys, 2)
    assert np.array_equal(out[0], ref.ct_to_words(up_ref))
    # decrypted control action vs the plaintext computation (CKKS error; tolerance 1e-7 on O(1) values): the ENGINE's
    # ciphertext goes through the reference's he_dec / he_dcd
    dec = ref.decode(ref.dec(ref.ct_from_words(up.download()[0], up.l, nu=up.nu, B=up.B), sk))
    plain = vals["uhat"] - (Ka @ (vals["xhat"] - vals["xr"]) + Kb @ (vals["uhat"] - vals["ur"]))
    assert np.abs(dec - plain).max() < 1e-7


def test_gemv_fused_equals_stepwise(cfgA, monkeypatch):
    """The memoised / NTT-domain-accumulated he_gemv schedule against the call-by-call reference schedule on a
    batch of random ciphertexts and a dense matrix (size-independent property: both must give the same words)."""
    from hectr_b200 import Engine
    eng, keys = cfgA["eng"], cfgA["keys"]
    rng = np.random.default_rng(43)
    S, l = 16, 2
    a = random_ct_words(rng, 5, eng.words(l), eng.n, eng.logq(l))
    A = rng.standard_normal((S, S)) + 1j * rng.standard_normal((S, S))
    fused = eng.he_gemv(A, eng.ct(a, l), keys).download()
    monkeypatch.setenv("HB_GEMV_STEPWISE", "1")
    eng2 = Engine(*A_PARAMS)
    try:
        rk_arr = np.stack([cfgA["ref"].evk_to_array(cfgA["rk"][i]) for i in range(S)])
        keys2 = eng2.keys(rk_arr)
        step = eng2.he_gemv(A, eng2.ct(a, l), keys2).download()
        keys2.free()
    finally:
        eng2.close()
    assert np.array_equal(fused, step)


def test_eval_host_pipeline_matches_device(cfgA, monkeypatch):
    """hb_ctr_hempc_eval_host cuts the batch into chunks that flow through upload / evaluate / download streams: ragged
    chunks (7 instances in chunks of 3) and a one-chunk batch must give the words of the device-resident call, and the
    hoisted baby rotations (one forward transform for n1 key switches) the words of one key switch per rotation."""
    from hectr_b200 import Engine
    eng, keys = cfgA["eng"], cfgA["keys"]
    rng = np.random.default_rng(47)
    S, l = 16, 2
    Ka = np.zeros((S, S), dtype=np.complex128)
    Kb = np.zeros((S, S), dtype=np.complex128)
    Ka[:8, :3] = rng.standard_normal((8, 3))
    Kb[:8, :2] = rng.standard_normal((8, 2))
    w = [random_ct_words(rng, 7, eng.words(l), eng.n, eng.logq(l)) for _ in range(4)]
    want = eng.ctr_hempc_eval(Ka, Kb, *[eng.ct(x, l) for x in w], keys).download()
    monkeypatch.setenv("HB_E2E_CHUNK", "3")
    assert np.array_equal(eng.ctr_hempc_eval_host(Ka, Kb, *w, keys, l), want)
    monkeypatch.setenv("HB_E2E_CHUNK", "64")
    assert np.array_equal(eng.ctr_hempc_eval_host(Ka, Kb, *w, keys, l), want)
    monkeypatch.setenv("HB_GEMV_NOHOIST", "1")
    eng2 = Engine(*A_PARAMS)
    try:
        rk_arr = np.stack([cfgA["ref"].evk_to_array(cfgA["rk"][i]) for i in range(S)])
        keys2 = eng2.keys(rk_arr)
        assert np.array_equal(eng2.ctr_hempc_eval(Ka, Kb, *[eng2.ct(x, l) for x in w], keys2).download(), want)
        keys2.free()
    finally:
        eng2.close()


def test_mul_relin(cfgA):
    ref, eng, pk, sk = cfgA["ref"], cfgA["eng"], cfgA["pk"], cfgA["sk"]
    rlk = ref.genrlk(sk)
    ekeys = eng.keys(ref.evk_to_array(rlk))
    rng = np.random.default_rng(51)
    for l in (2, 1):
        a = random_ct_words(rng, 2, eng.words(l), eng.n, eng.logq(l))
        b = random_ct_words(rng, 2, eng.words(l), eng.n, eng.logq(l))
        got = eng.he_mul(eng.ct(a, l), eng.ct(b, l), ekeys).download()
        for bi in range(2):
            want = ref.mul(_ref_ct(ref, a[bi], l), _ref_ct(ref, b[bi], l), rlk)
            assert np.array_equal(got[bi], ref.ct_to_words(want)), (l, bi)
    ekeys.free()


def test_enc_pk_and_dec_on_device(cfgA):
    """he_enc_pk / he_dec on the device with the REFERENCE's samples (sampling order defines every
    ciphertext, so the host draws v, e0, e1 in he_enc_pk's order: sample_zo, sample_error, sample_error)."""
    from conftest import ints_to_words, words_to_ints
    ref, eng, pk, sk = cfgA["ref"], cfgA["eng"], cfgA["pk"], cfgA["sk"]
    rng = np.random.default_rng(61)
    S, L, n = 16, ref.L, ref.n
    W, k = eng.words(L), eng.logq(L)
    q = 1 << k
    B = 3
    msgs = rng.standard_normal((B, S)) + 1j * rng.standard_normal((B, S))
    pts = [ref.encode(m) for m in msgs]
    pt_words = np.stack([ref.poly_to_words(p.m, W) for p in pts])
    v = np.stack([ref.sample_zo() for _ in range(B)])          # the draws interleave per instance in the reference;
    e0 = np.stack([ref.sample_error() for _ in range(B)])      # here any fixed order does: the arithmetic is what is compared
    e1 = np.stack([ref.sample_error() for _ in range(B)])
    assert set(np.unique(v)) <= {-1, 0, 1}
    dpk = eng.pk(ref.poly_to_words(pk.p0, W), ref.poly_to_words(pk.p1, W))
    got = eng.he_enc_pk(pt_words, dpk, v, e0, e1)
    assert got.l == L
    gw = got.download()
    smod = lambda x: ((x + q // 2) % q) - q // 2
    for b in range(B):
        vw = v[b].view(np.uint64)[None]
        prod0 = words_to_ints(ref.poly_mul_words(pk.p0, vw, ref.dim, k, W))          # poly_mul(pk.p0, v, hectx.dim, q)
        prod1 = words_to_ints(ref.poly_mul_words(pk.p1, vw, ref.dim, k, W))
        m = words_to_ints(pt_words[b])
        c0 = [smod(prod0[i] + m[i] + int(e0[b, i])) for i in range(n)]
        c1 = [smod(prod1[i] + int(e1[b, i])) for i in range(n)]
        assert np.array_equal(gw[b, 0], ints_to_words(c0, W)), b
        assert np.array_equal(gw[b, 1], ints_to_words(c1, W)), b
    # he_dec: device against the reference on (a) the device-encrypted batch and (b) a reference-encrypted ciphertext
    dsk = eng.sk(ref.poly_to_words(sk, 1)[0].view(np.int64))
    dec = eng.he_dec(got, dsk)
    for b in range(B):
        want = ref.dec(ref.ct_from_words(gw[b], L, nu=got.nu, B=got.B), sk)
        assert np.array_equal(dec[b], ref.poly_to_words(want.m, W)), b
        assert np.abs(ref.decode(want) - msgs[b]).max() < 1e-9          # and it decrypts to the message
    cref = ref.enc_pk(pts[0], pk)
    for l in (L, L - 1):
        wl = eng.words(l)
        assert np.array_equal(eng.he_dec(eng.ct(ref.ct_to_words(cref), l), dsk)[0], ref.poly_to_words(ref.dec(cref, sk).m, wl)), l
        ref.moddown(cref)


def test_config_C_full_chain_ring(ref_available):
    """BASELINE configs[2]: N = 2^15, q = 2^881 (the security-table maximum, precomp.c:63), Delta = 2^50: L = 17, key
    switches over 47 limbs, 14-word coefficients -- the multiword reconstruction and key-switch tail, clusters of 4 CTAs
    per limb.  he_mul + relinearise, he_rs and he_conj at the top level and at level 1 against the compiled reference."""
    from oracle.ref import Ref
    from hectr_b200 import Engine
    P = (15, 881, 16, 50)
    ref = Ref.get(*P)
    eng = Engine(*P)
    try:
        i = eng.info
        assert (i.L, i.dim, i.dimevk, i.dimub) == (ref.L, ref.dim, ref.dimevk, ref.dimub) == (17, 16, 47, 61)
        pk, sk = ref.keypair()
        rlk, ck = ref.genrlk(sk), ref.genck(sk)
        erlk, eck = eng.keys(ref.evk_to_array(rlk)), eng.keys(ref.evk_to_array(ck))
        rng = np.random.default_rng(71)
        for l in (ref.L, 1):
            W, k = eng.words(l), eng.logq(l)
            w = random_ct_words(rng, 2, W, eng.n, k)
            a, b = _ref_ct(ref, w[0], l), _ref_ct(ref, w[1], l)
            want = ref.mul(a, b, rlk)
            got = eng.he_mul(eng.ct(w[0], l), eng.ct(w[1], l), erlk)
            assert np.array_equal(got.download()[0], ref.ct_to_words(want)), ("he_mul", l)
            ref.rs(want)
            eng.he_rs(got)
            assert got.l == want.l == l - 1
            assert np.array_equal(got.download()[0], ref.ct_to_words(want)), ("he_rs", l)
            wantc = ref.conj(_ref_ct(ref, w[0], l), ck)
            gotc = eng.he_conj(eng.ct(w[0], l), eck)
            assert np.array_equal(gotc.download()[0], ref.ct_to_words(wantc)), ("he_conj", l)
        erlk.free()
        eck.free()
    finally:
        eng.close()


def test_config_D_gemv_32_slots(ref_available):
    """BASELINE configs[3] (horizon 10 -> slots = 32): n1 = 8 baby steps (two hoisted launches of four rotations), n2 = 4
    giant steps, 32 rotation keys, dense complex matrix; he_gemv against the compiled reference."""
    from oracle.ref import Ref
    from hectr_b200 import Engine
    P = (12, 109, 32, 50)
    ref = Ref.get(*P)
    eng = Engine(*P)
    try:
        pk, sk = ref.keypair()
        rk = ref.genrk(sk)
        keys = eng.keys(np.stack([ref.evk_to_array(rk[i]) for i in range(32)]))
        rng = np.random.default_rng(73)
        A = rng.standard_normal((32, 32)) + 1j * rng.standard_normal((32, 32))
        v = rng.standard_normal(32) + 1j * rng.standard_normal(32)
        ct = ref.enc_pk(ref.encode(v), pk)
        want = ref.gemv(A, ct, rk)
        got = eng.he_gemv(A, eng.ct(ref.ct_to_words(ct), ref.L, nu=ct.nu, B=ct.B), keys)
        assert got.l == want.l
        assert np.array_equal(got.download()[0], ref.ct_to_words(want))
        dec = ref.decode(ref.dec(want, sk))
        assert np.abs(dec - A @ v).max() < 1e-6
        keys.free()
    finally:
        eng.close()


def test_fast_tail_equals_exact_tail(cfgA, monkeypatch):
    """The fixed-point key-switch tail (fractions from the high halves, exact fallback near a decision boundary) against
    the multiword tail on 2 x 256 x 4096 coefficients per level: enough for a few dozen threads to land inside the margin
    (about 1e-5 per coefficient), so both branches and the hand-off between them are exercised."""
    from hectr_b200 import Engine
    eng, keys = cfgA["eng"], cfgA["keys"]
    rng = np.random.default_rng(83)
    S = 16
    monkeypatch.setenv("HB_KS_EXACT", "1")
    eng2 = Engine(*A_PARAMS)
    try:
        rk_arr = np.stack([cfgA["ref"].evk_to_array(cfgA["rk"][i]) for i in range(S)])
        keys2 = eng2.keys(rk_arr)
        for l in (2, 1):
            a = random_ct_words(rng, 256, eng.words(l), eng.n, eng.logq(l))
            fast = eng.he_rot(eng.ct(a, l), 5, keys).download()
            exact = eng2.he_rot(eng2.ct(a, l), 5, keys2).download()
            assert np.array_equal(fast, exact), l
        keys2.free()
    finally:
        eng2.close()


@pytest.mark.parametrize("logn,logq,logDelta,slots,level", [(12, 109, 50, 16, 2), (12, 109, 50, 16, 1), (12, 109, 50, 64, 2),
                                                             (11, 50, 24, 32, 2), (13, 218, 50, 16, 4)])
def test_pt30_basis_equals_59bit_basis(logn, logq, logDelta, slots, level, monkeypatch):
    """he_gemv's plaintext products run on the engine's internal 30-bit basis when the exact product fits P_D
    (hb_tables32.hpp); HB_NO_PT30 keeps them on the reference's 59-bit primes.  Both must give the same words: 128
    random ciphertexts at full size, dense random matrices (every diagonal non-zero), synthetic keys; every ring size the
    30-bit kernels are built for (their pass structure and bank-conflict lane swap depend on N)."""
    from hectr_b200 import Engine
    params = (logn, logq, slots, logDelta)       # q at most the reference's bound for the ring (precomp.c:338-350)
    rng = np.random.default_rng(97 + slots + level + logn)
    eng = Engine(*params)
    monkeypatch.setenv("HB_NO_PT30", "1")
    eng59 = Engine(*params)
    monkeypatch.delenv("HB_NO_PT30")
    try:
        primes = eng.primes()
        keys = np.empty((slots, 2, eng.info.dimevk, eng.n), dtype=np.uint64)
        for d in range(eng.info.dimevk):
            keys[:, :, d, :] = rng.integers(0, int(primes[d]), size=(slots, 2, eng.n), dtype=np.uint64)
        rk, rk59 = eng.keys(keys), eng59.keys(keys)
        K = rng.standard_normal((slots, slots)) + 1j * rng.standard_normal((slots, slots))
        a = random_ct_words(rng, 128, eng.words(level), eng.n, eng.logq(level))
        names = []
        for e in (eng, eng59):
            e.lib.hb_prof_reset(e.ctx)
            e.lib.hb_prof_enable(e.ctx, 1)
        got = eng.he_gemv(K, eng.ct(a, level), rk).download()
        want = eng59.he_gemv(K, eng59.ct(a, level), rk59).download()
        for e in (eng, eng59):
            e.lib.hb_prof_enable(e.ctx, 0)
            need = e.lib.hb_prof_report(e.ctx, None, 0)
            buf = C.create_string_buffer(need + 1)
            e.lib.hb_prof_report(e.ctx, buf, need + 1)
            names.append(set(json.loads(buf.value.decode())))
        assert any(n.startswith("k32_mulacc_intt") for n in names[0]), names[0]          # the 30-bit kernels really ran
        assert not any(n.startswith("k32_") for n in names[1]) and "k_mulacc_intt" in names[1], names[1]
        assert np.array_equal(got, want)
        rk.free()
        rk59.free()
    finally:
        eng.close()
        eng59.close()
