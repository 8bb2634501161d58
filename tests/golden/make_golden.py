#!/usr/bin/env python
"""Generates the committed golden vectors from the compiled, UNMODIFIED reference (oracle/_ref, built by
oracle/build_ref.sh from /root/reference).  Run where /root/reference exists:

    python tests/golden/make_golden.py

Outputs (small, committed):
  tests/golden/small_logn7.npz   -- every input and output of the evaluation path at logn=7, q=2^61, Delta=2^50,
                                    slots=16 (the ring of GPQHE/tests/polymul.gp): tables, keys, NTT pairs, and the
                                    result words of he_add/sub/neg/rs/moddown/mulpt/rot/conj/mul/gemv and of the
                                    ctr_hempc ciphertext sequence (src/hempc.c:253-266).  ~250 KiB.
  tests/golden/digests.json      -- SHA-256 of the reference's result words at the cstr-hempc parameters (logn=12,
                                    q=2^109) and at logn=13 / q=2^218 for inputs that are a pure function of a numpy
                                    seed (PCG64), so the GPU box can check the CUDA path without the vectors themselves.

The reference's RNG is the deterministic SUPERCOP stream (GPQHE/src/rng.c:32-78): keys are a pure function of the
call order below.  Everything the digests need besides the seed (the keys) is regenerated on the GPU box from
oracle/_ref when present; when it is not, only small_logn7.npz is used.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.ref import Ref  # noqa: E402
from conftest import random_ct_words  # noqa: E402


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()


def synth_keys(rng, primes, nkeys, dimevk, n):
    """Uniform key limbs below each prime: what the arithmetic sees; a pure function of the numpy seed."""
    keys = np.empty((nkeys, 2, dimevk, n), dtype=np.uint64)
    for d in range(dimevk):
        keys[:, :, d, :] = rng.integers(0, int(primes[d]), size=(nkeys, 2, n), dtype=np.uint64)
    return keys


def evk_from_array(ref, arr):
    """Write uint64[2][dimevk][N] into a freshly allocated reference he_evk_t."""
    import ctypes as C
    from oracle.ref import HeEvk
    evk = HeEvk()
    ref.lib.he_alloc_evk(C.byref(evk))
    shape = (ref.dimevk, ref.n)
    np.ctypeslib.as_array(evk.p0.coeffs, shape=shape)[:] = arr[0]
    np.ctypeslib.as_array(evk.p1.coeffs, shape=shape)[:] = arr[1]
    return evk


def evk_array_from(ref, arrs):
    import ctypes as C
    from oracle.ref import HeEvk
    rk = (HeEvk * len(arrs))()
    shape = (ref.dimevk, ref.n)
    for i, a in enumerate(arrs):
        ref.lib.he_alloc_evk(C.byref(rk[i]))
        np.ctypeslib.as_array(rk[i].p0.coeffs, shape=shape)[:] = a[0]
        np.ctypeslib.as_array(rk[i].p1.coeffs, shape=shape)[:] = a[1]
    return rk


def hempc_matrices(rng, S):
    Ka = np.zeros((S, S), dtype=np.complex128)
    Kb = np.zeros((S, S), dtype=np.complex128)
    Ka[:8, :3] = rng.standard_normal((8, 3))
    Kb[:8, :2] = rng.standard_normal((8, 2))
    return Ka, Kb


def diag_points(ref, A):
    """The slots encoded diagonals he_gemv builds (he-algo.c:64-72), as words[slots][W_top][N], order i*n1+j."""
    import math
    S = ref.slots
    n1 = int(math.sqrt(S))
    if S != n1 * n1:
        n1 = int(math.sqrt(2 * S))
    n2 = S // n1
    Wt = ref.words_per_level(ref.L)
    out = np.zeros((S, Wt, ref.n), dtype=np.uint64)
    for i in range(n2):
        shift = i * n1
        for j in range(n1):
            idx = shift + j
            diag = np.array([A[r % S, (idx + r) % S] for r in range(S)])
            rd = np.array([diag[(r - shift) % S] for r in range(S)])
            out[idx] = ref.poly_to_words(ref.encode(rd).m, Wt)
    return out


def small():
    P = (7, 61, 16, 50)
    ref = Ref.get(*P)
    rng = np.random.default_rng(20261104)
    n, S = ref.n, ref.slots
    out = {"params": np.array(P), "primes": np.array(ref.primes(), dtype=np.uint64),
           "derived": np.array([ref.L, ref.dim, ref.dimevk, ref.dimub])}
    out["zetas0"] = ref.zetas(0)
    out["zetas_inv0"] = ref.zetas(0, inv=True)
    # transforms: random limbs plus the adversarial ones of SURVEY 8(d) (zeros, p-1, and a forced "sum == p")
    p0 = ref.primes()[0]
    x = rng.integers(0, p0, size=(4, n), dtype=np.uint64)
    x[1] = 0
    x[2] = p0 - 1
    x[3, : n // 2] = p0
    x[3, n // 2:] = 0
    out["ntt_in"] = x
    out["ntt_out"] = np.stack([ref.ntt(v, 0) for v in x])          # values in [0, p]
    y = rng.integers(0, p0, size=(2, n), dtype=np.uint64)
    out["intt_in"] = y
    out["intt_out"] = np.stack([ref.ntt(v, 0, inverse=True) for v in y])
    # keys from the reference's deterministic stream
    pk, sk = ref.keypair()
    rk = ref.genrk(sk)
    rlk = ref.genrlk(sk)
    ck = ref.genck(sk)
    out["rk"] = np.stack([ref.evk_to_array(rk[i]) for i in range(S)])
    out["rlk"] = ref.evk_to_array(rlk)
    out["ck"] = ref.evk_to_array(ck)
    l = ref.L
    W = ref.words_per_level(l)
    msgs = rng.standard_normal((4, S)) + 1j * rng.standard_normal((4, S))
    cts = [ref.enc_pk(ref.encode(m), pk) for m in msgs]
    out["msgs"] = msgs
    out["ct_in"] = np.stack([ref.ct_to_words(c) for c in cts])          # [4][2][W][N]
    a, b = cts[0], cts[1]
    out["add"] = ref.ct_to_words(ref.add(a, b))
    out["sub"] = ref.ct_to_words(ref.sub(a, b))
    out["neg"] = ref.ct_to_words(ref.neg(ref.copy_ct(a)))
    out["rs"] = ref.ct_to_words(ref.rs(ref.copy_ct(a)))
    out["moddown"] = ref.ct_to_words(ref.moddown(ref.copy_ct(a)))
    ptm = rng.standard_normal(S) + 1j * rng.standard_normal(S)
    pt = ref.encode(ptm)
    out["pt_words"] = ref.poly_to_words(pt.m, W)
    out["mulpt"] = ref.ct_to_words(ref.mulpt(a, pt))
    out["rot"] = np.stack([ref.ct_to_words(ref.rot(ref.copy_ct(a), r, rk)) for r in range(S)])
    out["conj"] = ref.ct_to_words(ref.conj(ref.copy_ct(a), ck))
    out["mul"] = ref.ct_to_words(ref.mul(a, b, rlk))
    Ka, Kb = hempc_matrices(rng, S)
    out["Ka"], out["Kb"] = Ka, Kb
    out["Ka_pts"], out["Kb_pts"] = diag_points(ref, Ka), diag_points(ref, Kb)
    g = ref.gemv(Ka, a, rk)
    out["gemv"] = ref.ct_to_words(g)
    xhat, xr, uhat, ur = cts
    du = ref.neg(ref.add(ref.gemv(Ka, ref.sub(xhat, xr), rk), ref.gemv(Kb, ref.sub(uhat, ur), rk)))
    up = ref.add(ref.moddown(ref.copy_ct(uhat)), du)
    out["hempc"] = ref.ct_to_words(up)
    dec = ref.decode(ref.dec(up, sk))
    out["hempc_decoded"] = dec
    np.savez_compressed(os.path.join(HERE, "small_logn7.npz"), **out)
    print("small_logn7.npz:", os.path.getsize(os.path.join(HERE, "small_logn7.npz")), "bytes")


def digests():
    res = {}
    for tag, P, seed in (("A", (12, 109, 16, 50), 4096), ("B13", (13, 218, 16, 50), 8192)):
        ref = Ref.get(*P)
        rng = np.random.default_rng(seed)
        n, S, l = ref.n, ref.slots, ref.L
        W, k = ref.words_per_level(l), ref.q_level(l).bit_length() - 1
        keys = synth_keys(rng, ref.primes(), S + 1, ref.dimevk, n)       # rk[0..S), then rlk
        rk = evk_array_from(ref, keys[:S])
        rlk = evk_from_array(ref, keys[S])
        w = random_ct_words(rng, 4, W, n, k)
        cts = [ref.ct_from_words(w[i], l) for i in range(4)]
        d = {"params": list(P), "seed": seed, "inputs": sha(w), "keys": sha(keys)}
        d["sub"] = sha(ref.ct_to_words(ref.sub(cts[0], cts[1])))
        d["rot3"] = sha(ref.ct_to_words(ref.rot(ref.copy_ct(cts[0]), 3, rk)))
        d["rs"] = sha(ref.ct_to_words(ref.rs(ref.copy_ct(cts[0]))))
        d["mul"] = sha(ref.ct_to_words(ref.mul(cts[0], cts[1], rlk)))
        if tag == "A":
            Ka, Kb = hempc_matrices(rng, S)
            d["gemv"] = sha(ref.ct_to_words(ref.gemv(Ka, cts[0], rk)))
            xhat, xr, uhat, ur = cts
            du = ref.neg(ref.add(ref.gemv(Ka, ref.sub(xhat, xr), rk), ref.gemv(Kb, ref.sub(uhat, ur), rk)))
            d["hempc"] = sha(ref.ct_to_words(ref.add(ref.moddown(ref.copy_ct(uhat)), du)))
        res[tag] = d
        print(tag, d)
    json.dump(res, open(os.path.join(HERE, "digests.json"), "w"), indent=1)


if __name__ == "__main__":
    small()
    digests()
