"""Parity at the shapes bench.py publishes (VERDICT r01 "next" item 1): the 4096-instance launches of config E and the chunked
host pipeline, config D at 64 and 128 slots, the NTT with 1 / 8 / 24 limbs per launch at N = 2^13..2^16, and the accepted
parameter corner where he_mulpt needs a single limb.  Everything bit-exact against the compiled, unmodified reference."""
import ctypes as C
import json

import numpy as np
import pytest

from conftest import random_ct_words

pytestmark = pytest.mark.gpu

A_PARAMS = (12, 109, 16, 50)


def hempc_gains(rng, S=16, nx=3, nu=2, horizon=4):
    """Ka (mN x n) and Kb (mN x m) in a slots x slots frame, as ctr_hempc builds them (src/hempc.c:232-238)."""
    Ka = np.zeros((S, S), dtype=np.complex128)
    Kb = np.zeros((S, S), dtype=np.complex128)
    Ka[: nu * horizon, :nx] = rng.standard_normal((nu * horizon, nx))
    Kb[: nu * horizon, :nu] = rng.standard_normal((nu * horizon, nu))
    return Ka, Kb


def ref_hempc(ref, rk, Ka, Kb, c):
    """src/hempc.c:253-266 through the reference's own calls."""
    xd, ud = ref.sub(c["xhat"], c["xr"]), ref.sub(c["uhat"], c["ur"])
    du = ref.neg(ref.add(ref.gemv(Ka, xd, rk), ref.gemv(Kb, ud, rk)))
    return ref.add(ref.moddown(ref.copy_ct(c["uhat"])), du)


@pytest.fixture(scope="module")
def cfgA(ref_available):
    from oracle.ref import Ref
    from hectr_b200 import Engine
    ref = Ref.get(*A_PARAMS)
    eng = Engine(*A_PARAMS)
    pk, sk = ref.keypair()
    rk = ref.genrk(sk)
    keys = eng.keys(np.stack([ref.evk_to_array(rk[i]) for i in range(ref.slots)]))
    yield dict(ref=ref, eng=eng, pk=pk, sk=sk, rk=rk, keys=keys)
    keys.free()
    eng.close()


PROBES = (0, 1, 15, 16, 511, 512, 4095)


def test_config_E_4096_instances(cfgA, monkeypatch):
    """The launches bench.py times: hb_ctr_hempc_eval on 4096 instances, hb_ctr_hempc_eval_host on 4096 (512-instance chunks)
    and on a 512-instance shard (the chunking an 8-GPU rank sees).  Seven distinct reference-encrypted controller states sit
    at the probe indices and are compared word for word with the oracle; every other instance is a copy of one of the seven
    and must reproduce its words (grid, group and stride arithmetic at full size)."""
    ref, eng, keys, pk, sk, rk = (cfgA[k] for k in ("ref", "eng", "keys", "pk", "sk", "rk"))
    rng = np.random.default_rng(4096)
    S, B, l = 16, 4096, 2
    Ka, Kb = hempc_gains(rng)
    names = ("xhat", "xr", "uhat", "ur")
    states, cts = [], []
    for _ in PROBES:
        vals, c = {}, {}
        for nm in names:
            v = np.zeros(S, dtype=np.complex128)
            v[: 3 if nm[0] == "x" else 2] = rng.standard_normal(3 if nm[0] == "x" else 2)
            vals[nm] = v
            c[nm] = ref.enc_pk(ref.encode(v), pk)
        states.append(vals)
        cts.append(c)
    want = [ref.ct_to_words(ref_hempc(ref, rk, Ka, Kb, c)) for c in cts]
    src = np.arange(B) % len(PROBES)                       # instance b carries distinct state src[b]
    src = (src * 5 + (np.arange(B) // 7)) % len(PROBES)    # scattered
    for s, b in enumerate(PROBES):
        src[b] = s
    W = eng.words(l)
    host = {}
    for nm in names:
        seven = np.stack([ref.ct_to_words(c[nm]) for c in cts])           # [7][2][W][N]
        host[nm] = np.ascontiguousarray(seven[src])
        assert host[nm].shape == (B, 2, W, eng.n)

    def check(out, label):
        for s, b in enumerate(PROBES):
            if b < out.shape[0]:
                assert np.array_equal(out[b], want[s]), (label, "probe", b)
        assert np.array_equal(out, np.stack(want)[src[: out.shape[0]]]), (label, "copies")

    dev = [eng.ct(host[nm], l, nu=cts[0][nm].nu, B=cts[0][nm].B) for nm in names]
    up = eng.ctr_hempc_eval(Ka, Kb, *dev, keys)
    check(up.download(), "device-resident 4096")
    # the engine's own result decrypts to the control action (reference he_dec / he_dcd on the ENGINE's ciphertext)
    out = up.download()
    for s, b in enumerate(PROBES[:3]):
        st = states[s]
        plain = st["uhat"] - (Ka @ (st["xhat"] - st["xr"]) + Kb @ (st["uhat"] - st["ur"]))
        dec = ref.decode(ref.dec(ref.ct_from_words(out[b], l - 1, nu=up.nu, B=up.B), sk))
        assert np.abs(dec - plain).max() < 1e-7, b
    up.free()
    for d in dev:
        d.free()
    check(eng.ctr_hempc_eval_host(Ka, Kb, *[host[nm] for nm in names], keys, l), "host pipeline 4096 (default chunks)")
    check(eng.ctr_hempc_eval_host(Ka, Kb, *[host[nm][:512] for nm in names], keys, l), "host pipeline 512 (8-GPU shard)")


# (the tests that share the module's config-A reference context come first: the reference keeps ONE global context and
# every Ref.get with other parameters re-initialises it)
def test_gemv_shortcuts_equal_the_plain_schedule(cfgA, monkeypatch):
    """Three exact shortcuts of he_gemv, all on by default and all bound-checked against the measured key size:
      * identically zero giant steps are skipped                              (HB_GEMV_NOSKIP=1 switches it off)
      * the giant-step key switches share one reconstruction                  (HB_GEMV_NOSUM=1;  k_ks_limb_sum / k_ks_finish_sum)
      * the key-switch limbs that do not divide P run on 30-bit primes        (HB_GEMV_NOHYB=1;  k32_ks_baby / k32_ks_limb_sum / k_ks_finish_*hyb)
    Engines with (all on), (hybrid off) and (everything off) must give the same words on 192 full-size random ciphertexts
    per level -- a few of the 3 M coefficients land inside the fixed-point margins and take the exact paths; one instance has
    c1 = 0, where EVERY decision about R sits on a boundary -- for a dense matrix and for the hempc block matrices; the launch
    profile names which kernels ran."""
    from hectr_b200 import Engine
    eng, ref = cfgA["eng"], cfgA["ref"]
    rk_arr = np.stack([ref.evk_to_array(cfgA["rk"][i]) for i in range(16)])
    rng = np.random.default_rng(811)
    Ka, Kb = hempc_gains(rng)
    Ad = rng.standard_normal((16, 16)) + 1j * rng.standard_normal((16, 16))

    def banded(diags):
        """only the generalised diagonals `diags` are non-zero (zrotdiag, he-algo.c:29-43: entry r of diagonal d is A[r, r + d])"""
        A = np.zeros((16, 16), dtype=np.complex128)
        for d in diags:
            for r in range(16):
                A[r, (r + d) % 16] = rng.standard_normal() + 1j * rng.standard_normal()
        return A

    M1, M2 = banded(range(0, 4)), banded(list(range(0, 4)) + list(range(8, 12)))     # one / two giant steps of four evaluated
    monkeypatch.setenv("HB_GEMV_NOHYB", "1")
    sum59 = Engine(*A_PARAMS)
    monkeypatch.setenv("HB_GEMV_NOSUM", "1")
    monkeypatch.setenv("HB_GEMV_NOSKIP", "1")
    plain = Engine(*A_PARAMS)
    for v in ("HB_GEMV_NOSUM", "HB_GEMV_NOSKIP", "HB_GEMV_NOHYB"):
        monkeypatch.delenv(v)

    def names_of(e):
        need = e.lib.hb_prof_report(e.ctx, None, 0)
        buf = C.create_string_buffer(need + 1)
        e.lib.hb_prof_report(e.ctx, buf, need + 1)
        return json.loads(buf.value.decode())

    try:
        keys_s, keys_p = sum59.keys(rk_arr), plain.keys(rk_arr)
        engines = ((eng, cfgA["keys"]), (sum59, keys_s), (plain, keys_p))
        for l in (2, 1):
            a = random_ct_words(rng, 192, eng.words(l), eng.n, eng.logq(l))
            a[0, 1] = 0                                     # c1 = 0: every R is exactly 0
            for M in (Ad, Ka, Kb, M1, M2):
                outs, profs = [], []
                for e, kk in engines:
                    e.lib.hb_prof_reset(e.ctx)
                    e.lib.hb_prof_enable(e.ctx, 1)
                    outs.append(e.he_gemv(M, e.ct(a, l), kk))
                    profs.append(names_of(e))
                    e.lib.hb_prof_enable(e.ctx, 0)
                ph, ps, pp = profs
                assert {"k32_ks_baby", "k_ks_finish_hyb", "k32_ks_limb_sum", "k_ks_finish_sum_hyb"} <= set(ph), sorted(ph)
                assert {"k_ks_limb_sum", "k_ks_finish_sum"} <= set(ps) and not any("hyb" in n or n.startswith("k32_ks") for n in ps), sorted(ps)
                assert "k_ct_sum_rs" in pp and not any("sum" in n and n != "k_ct_sum_rs" for n in pp) and not any("hyb" in n for n in pp), sorted(pp)
                want = outs[2].download()
                for o in outs[:2]:
                    assert o.l == outs[2].l and o.nu == outs[2].nu and o.B == outs[2].B
                    assert np.array_equal(o.download(), want), l
                for o in outs:
                    o.free()
        keys_s.free()
        keys_p.free()
    finally:
        sum59.close()
        plain.close()


def test_paired_matvecs_equal_two_passes(cfgA, monkeypatch):
    """hb_ctr_hempc_eval runs Ka (xhat - xr) and Kb (uhat - ur) as ONE pass over a batch of 2 x B instances (only the diagonal
    products know which matrix an instance belongs to); HB_HEMPC_NOPAIR=1 keeps two passes.  Same words on 96 random
    controller states, with the hempc block matrices and with dense ones (different sets of non-zero giant steps)."""
    from hectr_b200 import Engine
    eng, ref = cfgA["eng"], cfgA["ref"]
    rk_arr = np.stack([ref.evk_to_array(cfgA["rk"][i]) for i in range(16)])
    rng = np.random.default_rng(911)
    monkeypatch.setenv("HB_HEMPC_NOPAIR", "1")
    two = Engine(*A_PARAMS)
    monkeypatch.delenv("HB_HEMPC_NOPAIR")
    try:
        keys2 = two.keys(rk_arr)
        l = 2
        w = [random_ct_words(rng, 96, eng.words(l), eng.n, eng.logq(l)) for _ in range(4)]
        Ka, Kb = hempc_gains(rng)
        dense = rng.standard_normal((16, 16)) + 1j * rng.standard_normal((16, 16))
        for A, B in ((Ka, Kb), (dense, Kb), (Ka, dense)):
            got = eng.ctr_hempc_eval(A, B, *[eng.ct(x, l) for x in w], cfgA["keys"])
            want = two.ctr_hempc_eval(A, B, *[two.ct(x, l) for x in w], keys2)
            assert got.l == want.l and got.nu == want.nu and got.B == want.B
            assert np.array_equal(got.download(), want.download())
        keys2.free()
    finally:
        two.close()


def test_streaming_host_interface(cfgA):
    """hb_ctr_hempc_submit_host / hb_ctr_hempc_wait: four batches of different sizes and contents submitted back to back (two in
    flight, the third submission waits for the first), waited for out of order; every result must equal the device-resident
    evaluation of the same batch, and a larger batch after smaller ones regrows the persistent buffer sets."""
    eng, keys = cfgA["eng"], cfgA["keys"]
    rng = np.random.default_rng(1201)
    l = 2
    Ka, Kb = hempc_gains(rng)
    sizes = (5, 9, 3, 17)
    ins, outs, want = [], [], []
    for nb in sizes:
        w = [random_ct_words(rng, nb, eng.words(l), eng.n, eng.logq(l)) for _ in range(4)]
        ins.append(w)
        outs.append(np.zeros((nb, 2, eng.words(l - 1), eng.n), dtype=np.uint64))
        want.append(eng.ctr_hempc_eval(Ka, Kb, *[eng.ct(x, l) for x in w], keys).download())
    tickets = [eng.ctr_hempc_submit_host(Ka, Kb, *ins[i], keys, l, outs[i]) for i in range(len(sizes))]
    for i in (1, 0, 3, 2):
        eng.ctr_hempc_wait(tickets[i])
        assert np.array_equal(outs[i], want[i]), i


def test_synthetic_keys_keep_the_per_step_tails(cfgA):
    """Uniform residues are not a switching key: their lift reaches P_D, the products wrap, and the sum of the wrapped
    reconstructions is not the reconstruction of the sum.  hb_evk_upload records the bound of every key (k_lift_bits) and
    he_gemv keeps one tail per giant step for such keys."""
    eng = cfgA["eng"]
    rng = np.random.default_rng(5)
    primes = eng.primes()
    fake = np.empty((16, 2, eng.info.dimevk, eng.n), dtype=np.uint64)
    for d in range(eng.info.dimevk):
        fake[:, :, d, :] = rng.integers(0, int(primes[d]), size=(16, 2, eng.n), dtype=np.uint64)
    keys = eng.keys(fake)
    A = rng.standard_normal((16, 16)) + 1j * rng.standard_normal((16, 16))
    a = random_ct_words(rng, 2, eng.words(2), eng.n, eng.logq(2))
    eng.lib.hb_prof_reset(eng.ctx)
    eng.lib.hb_prof_enable(eng.ctx, 1)
    eng.he_gemv(A, eng.ct(a, 2), keys).free()
    need = eng.lib.hb_prof_report(eng.ctx, None, 0)
    buf = C.create_string_buffer(need + 1)
    eng.lib.hb_prof_report(eng.ctx, buf, need + 1)
    eng.lib.hb_prof_enable(eng.ctx, 0)
    names = json.loads(buf.value.decode())
    assert "k_ks_finish_sum" not in names and "k_ct_sum_rs" in names, sorted(names)
    keys.free()


@pytest.mark.parametrize("slots,n1,n2", [(64, 8, 8), (128, 16, 8)])
def test_config_D_gemv_64_128_slots(ref_available, slots, n1, n2):
    """BASELINE configs[3], horizon 16-50: slots = 1 << (32 - clz(nu horizon)) = 64 / 128 (src/ctr.c:511), he_gemv with
    n1 = 8 / 16 baby steps (two / four hoisted launches) and n2 = 8 giant steps, `slots` rotation keys, dense complex
    matrix, against he_gemv of the compiled reference (GPQHE/src/he-algo.c:47-93)."""
    from oracle.ref import Ref
    from hectr_b200 import Engine
    P = (12, 109, slots, 50)
    ref = Ref.get(*P)
    eng = Engine(*P)
    try:
        pk, sk = ref.keypair()
        rk = ref.genrk(sk)
        keys = eng.keys(np.stack([ref.evk_to_array(rk[i]) for i in range(slots)]))
        rng = np.random.default_rng(slots)
        A = rng.standard_normal((slots, slots)) + 1j * rng.standard_normal((slots, slots))
        v = rng.standard_normal(slots) + 1j * rng.standard_normal(slots)
        ct = ref.enc_pk(ref.encode(v), pk)
        want = ref.gemv(A, ct, rk)
        w = ref.ct_to_words(ct)
        got = eng.he_gemv(A, eng.ct(np.stack([w, w, w]), ref.L, nu=ct.nu, B=ct.B), keys)
        assert got.l == want.l and got.nu == want.nu and got.B == want.B
        out = got.download()
        for b in range(3):
            assert np.array_equal(out[b], ref.ct_to_words(want)), b
        dec = ref.decode(ref.dec(ref.ct_from_words(out[0], got.l, nu=got.nu, B=got.B), sk))
        assert np.abs(dec - A @ v).max() < 1e-5
        # a sparse matrix with the hempc block structure at this size: identically zero giant steps are skipped
        K = np.zeros((slots, slots), dtype=np.complex128)
        K[: slots // 2, :3] = rng.standard_normal((slots // 2, 3))
        want2 = ref.gemv(K, ct, rk)
        got2 = eng.he_gemv(K, eng.ct(w, ref.L, nu=ct.nu, B=ct.B), keys)
        assert np.array_equal(got2.download()[0], ref.ct_to_words(want2))
        keys.free()
    finally:
        eng.close()


@pytest.mark.parametrize("logn", [13, 14, 15, 16])
@pytest.mark.parametrize("D", [1, 8, 24])
def test_ntt_limb_counts(ref_available, logn, D):
    """BASELINE configs[1]: batched forward / inverse NTT with 1, 8 and 24 limbs (primes) per launch at N = 2^13..2^16, the
    reference's prime chain for that ring, 64+ limbs per launch incl. the F8 adversarial limbs; bit-exact vs ntt / invntt
    (GPQHE/src/ntt.c:37-73) in the reference's own [0,p] representation."""
    from oracle.ref import Ref
    from hectr_b200 import Engine
    logq = {13: 218, 14: 438, 15: 881, 16: 1450}[logn]
    logq = min(logq, 1023)
    ref = Ref.get(logn, logq, 16, 50)
    eng = Engine(logn, logq, 16, 50)
    try:
        primes = ref.primes()
        D = min(D, len(primes))                            # the reference's chain for N = 2^13 at its largest q holds 16 primes
        n = eng.n
        rng = np.random.default_rng(logn * 100 + D)
        nl = max(64, 3 * D)
        nl -= nl % D
        limbs = np.empty((nl, n), dtype=np.uint64)
        for j in range(nl):
            limbs[j] = rng.integers(0, primes[j % D], size=n, dtype=np.uint64)
        limbs[0] = 0
        limbs[D % nl] = primes[D % nl % D] - 1
        limbs[-1, ::2] = primes[(nl - 1) % D]              # a[j] = p (F8): representable input of the reference's butterflies
        limbs[-1, 1::2] = 0
        got_f = eng.ntt(limbs, 0, D, strict=True)
        sample = sorted(set(list(range(0, nl, max(1, nl // 12))) + [0, D % nl, nl - 1]))
        for j in sample:
            assert np.array_equal(got_f[j], ref.ntt(limbs[j], j % D)), ("forward", j)
        inp = limbs.copy()
        inp[-1] = rng.integers(0, primes[(nl - 1) % D], size=n, dtype=np.uint64)
        got_i = eng.ntt(inp, 0, D, inverse=True)
        for j in sample:
            assert np.array_equal(got_i[j], ref.ntt(inp[j], j % D, inverse=True)), ("inverse", j)
    finally:
        eng.close()


@pytest.mark.parametrize("P", [(11, 46, 16, 18), (13, 107, 16, 21)])
def test_gemv_single_limb_plaintext_products(ref_available, P):
    """dim_mulpt == 1 (he-mult.c:168) happens at level 1 of short chains: the 30-bit basis then needs three limbs for
    a 60-bit P_1 -- more than twice the 59-bit limb count -- and used to overrun the shared workspaces.  (11, 2^46, 2^18) has a
    one-prime P (generic key-switch tail); (13, 2^107, 2^21) takes the fixed-point tails and the summed giant steps at N = 2^13."""
    from oracle.ref import Ref
    from hectr_b200 import Engine
    ref = Ref.get(*P)
    eng = Engine(*P)
    try:
        l = 1
        Delta = 2.0 ** P[3]
        assert eng.lib.hb_dim_mulpt(eng.ctx, l, Delta) == 1 and eng.logq(0) > 0
        pk, sk = ref.keypair()
        rk = ref.genrk(sk)
        keys = eng.keys(np.stack([ref.evk_to_array(rk[i]) for i in range(16)]))
        rng = np.random.default_rng(1154 + P[0])
        A = (rng.standard_normal((16, 16)) + 1j * rng.standard_normal((16, 16))) * 0.25
        w = random_ct_words(rng, 5, eng.words(l), eng.n, eng.logq(l))
        got = eng.he_gemv(A, eng.ct(w, l, nu=Delta, B=1.0), keys)
        out = got.download()
        for b in (0, 4):
            c = ref.ct_from_words(w[b], l, nu=Delta, B=1.0)
            want = ref.gemv(A, c, rk)
            assert got.l == want.l == l - 1
            assert np.array_equal(out[b], ref.ct_to_words(want)), b
        keys.free()
    finally:
        eng.close()


def test_config_C_full_level_chain(ref_available):
    """BASELINE configs[2] as SURVEY.md 8(d) words it: N = 2^15, q = 2^881, Delta = 2^50, reference-encrypted m1, m2 of
    GPQHE/tests/gpqhe.c:200-203, then at EVERY level 17..1: he_rot(a, 1), he_mul(a, b) + relinearise, he_rs, he_moddown(b),
    the ciphertext words compared after every operation.  The whole chain is pinned by tests/golden/config_c_chain.json
    (SHA-256 of the unmodified reference's ciphertexts, written by tests/golden/make_config_c.py); at levels 17, 9 and 1 the
    compiled reference is additionally run live on the engine's own operands.  Keys and inputs come from the reference on this
    host (deterministic stream); if they do not reproduce the committed digests, every level is checked live instead."""
    import hashlib
    import os
    from oracle.ref import Ref
    from hectr_b200 import Engine
    sys_path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_config_c", os.path.join(sys_path, "make_config_c.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    gold = json.load(open(os.path.join(sys_path, "config_c_chain.json")))
    P = tuple(gold["params"])
    ref = Ref.get(*P)
    eng = Engine(*P)
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()
    try:
        pk, sk, rlk, rk, a, b, m1, m2 = gen.setup(ref)
        wa, wb = ref.ct_to_words(a), ref.ct_to_words(b)
        same_inputs = (sha(wa) == gold["inputs"]["a"] and sha(wb) == gold["inputs"]["b"] and
                       sha(ref.evk_to_array(rlk)) == gold["inputs"]["rlk"] and sha(ref.evk_to_array(rk[1])) == gold["inputs"]["rk1"])
        live_levels = {ref.L, 9, 1} if same_inputs else set(range(1, ref.L + 1))
        erlk = eng.keys(ref.evk_to_array(rlk))
        erk = eng.keys(np.stack([ref.evk_to_array(rk[i]) for i in range(ref.slots)]))
        ea = eng.ct(wa, ref.L, nu=a.nu, B=a.B)
        eb = eng.ct(wb, ref.L, nu=b.nu, B=b.B)
        for l in range(ref.L, 0, -1):
            g = gold["levels"][str(l)]
            wa, wb = ea.download()[0], eb.download()[0]
            er = eng.he_rot(eng.he_copy_ct(ea), 1, erk)
            ep = eng.he_mul(ea, eb, erlk)
            w_rot, w_mul = er.download()[0], ep.download()[0]
            nu_mul, B_mul = ep.nu, ep.B
            eng.he_rs(ep)
            eng.he_moddown(eb)
            w_rs, w_md = ep.download()[0], eb.download()[0]
            assert ep.l == eb.l == l - 1
            if same_inputs:
                assert sha(w_rot) == g["rot"], ("he_rot", l)
                assert sha(w_mul) == g["mul"], ("he_mul", l)
                assert sha(w_rs) == g["rs"], ("he_rs", l)
                assert sha(w_md) == g["moddown"], ("he_moddown", l)
            if l in live_levels:
                ra, rb = ref.ct_from_words(wa, l, nu=ea.nu, B=ea.B), ref.ct_from_words(wb, l, nu=eb.nu, B=eb.B)
                rr = ref.rot(ref.copy_ct(ra), 1, rk)
                assert np.array_equal(w_rot, ref.ct_to_words(rr)), ("he_rot live", l)
                rp = ref.mul(ra, rb, rlk)
                assert np.array_equal(w_mul, ref.ct_to_words(rp)), ("he_mul live", l)
                assert rp.nu == nu_mul and abs(rp.B - B_mul) <= 1e-9 * abs(rp.B), ("he_mul metadata", l, rp.nu, nu_mul, rp.B, B_mul)
                ref.rs(rp)
                assert np.array_equal(w_rs, ref.ct_to_words(rp)), ("he_rs live", l)
                ref.moddown(rb)
                assert np.array_equal(w_md, ref.ct_to_words(rb)), ("he_moddown live", l)
                for c in (ra, rb, rr, rp):
                    ref.free_ct(c)
            er.free()
            ea.free()
            ea = ep
        # the level-0 product of the whole chain still decrypts to something finite: not a parity criterion, a smoke check
        erlk.free()
        erk.free()
    finally:
        eng.close()
