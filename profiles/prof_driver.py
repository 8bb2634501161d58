#!/usr/bin/env python
"""Small torch-free driver for ncu captures: launches the hot kernels on synthetic inputs through the C ABI.

    python profiles/prof_driver.py --case hempc --batch 4096     # one batched ctr_hempc evaluation (config A)
    python profiles/prof_driver.py --case ntt --logn 16          # batched forward + inverse NTT (config B)
    python profiles/prof_driver.py --case mul --logn 15 --logq 881 --batch 8    # ct x ct + relin + rescale + rot (config C)

Prints CUDA-event times so the same command doubles as the "plain run" that must exit 0 before ncu.
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hectr_b200 import Engine  # noqa: E402


def rand_ct(rng, batch, W, n, k):
    w = rng.integers(0, 1 << 64, size=(batch, 2, W, n), dtype=np.uint64)
    top = k - 64 * (W - 1)
    if top < 64:
        t = w[:, :, W - 1, :].view(np.int64)
        t <<= np.int64(64 - top)
        t >>= np.int64(64 - top)
    return w


def rand_keys(eng, rng, nkeys):
    """REAL switching keys (hb_he_genswk from a sampled secret, as bench.py makes them): the bound-checked shortcuts of he_gemv
    only apply to keys whose lift stays below P q_L / 2."""
    import bench
    secret = bench.make_secret(rng, eng.n)
    rots = bench.gemv_rotations(eng.slots) if nkeys == eng.slots else list(range(nkeys))
    return bench.make_rotation_keys(eng, rng, secret, rots)[:nkeys]


def prof_report(eng):
    need = eng.lib.hb_prof_report(eng.ctx, None, 0)
    buf = C.create_string_buffer(need + 1)
    eng.lib.hb_prof_report(eng.ctx, buf, need + 1)
    return json.loads(buf.value.decode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="hempc", choices=["hempc", "ntt", "mul"])
    ap.add_argument("--logn", type=int, default=12)
    ap.add_argument("--logq", type=int, default=109)
    ap.add_argument("--slots", type=int, default=16)
    ap.add_argument("--logDelta", type=int, default=50)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--limbs", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    rng = np.random.default_rng(1)
    eng = Engine(a.logn, a.logq, a.slots, a.logDelta, device=0)
    lib, ctx = eng.lib, eng.ctx
    n = eng.n
    if a.case == "ntt":
        D = min(8, eng.info.dimub)
        primes = eng.primes()
        host = np.empty((a.limbs, n), dtype=np.uint64)
        for j in range(a.limbs):
            host[j] = rng.integers(0, int(primes[j % D]), size=n, dtype=np.uint64)
        d_in = eng.to_device(host)
        d_out = eng.dev_alloc(host.nbytes)
        lib.hb_prof_enable(ctx, 1)
        for _ in range(a.reps + 1):
            lib.hb_ntt_forward(ctx, d_in, d_out, a.limbs, 0, D, 0)
            lib.hb_ntt_inverse(ctx, d_in, d_out, a.limbs, 0, D)
        rep = prof_report(eng)
        for kname, v in rep.items():
            ms = v["ms"] / v["launches"]
            print(f"{kname}: {ms:.4f} ms/launch  {a.limbs / ms * 1e3:.0f} limbs/s  {16.0 * n * a.limbs / ms / 1e6:.1f} GB/s")
    elif a.case == "hempc":
        l = eng.L
        W, k = eng.words(l), eng.logq(l)
        keys = eng.keys(rand_keys(eng, rng, a.slots))
        S = a.slots
        Ka = np.zeros((S, S), dtype=np.complex128)
        Kb = np.zeros((S, S), dtype=np.complex128)
        Ka[:8, :3] = rng.standard_normal((8, 3))
        Kb[:8, :2] = rng.standard_normal((8, 2))
        cts = []
        for _ in range(4):          # four DIFFERENT inputs: xhat - xr must not be the zero ciphertext
            base = rand_ct(rng, min(a.batch, 64), W, n, k)
            reps = (a.batch + base.shape[0] - 1) // base.shape[0]
            cts.append(eng.ct(np.concatenate([base] * reps)[: a.batch], l))
        out = eng.ct(batch=a.batch)
        eng.ctr_hempc_eval(Ka, Kb, *cts, keys, out=out)      # warm (builds the gemv plans)
        eng.sync()
        lib.hb_prof_enable(ctx, 1)
        lib.hb_timer_start(ctx)
        for _ in range(a.reps):
            eng.ctr_hempc_eval(Ka, Kb, *cts, keys, out=out)
        ms = lib.hb_timer_stop_ms(ctx) / a.reps
        print(f"ctr_hempc_eval batch={a.batch}: {ms:.2f} ms  -> {a.batch / ms * 1e3:.0f} iterations/s")
        rep = prof_report(eng)
        tot = sum(v["ms"] for v in rep.values())
        for kname, v in sorted(rep.items(), key=lambda kv: -kv[1]["ms"]):
            print(f"  {kname}: {v['launches']} launches, {v['ms'] / v['launches']:.4f} ms each, share {v['ms'] / tot:.3f}")
    else:
        l = eng.L
        W, k = eng.words(l), eng.logq(l)
        keys = eng.keys(rand_keys(eng, rng, 2))
        x = eng.ct(rand_ct(rng, a.batch, W, n, k), l)
        y = eng.ct(rand_ct(rng, a.batch, W, n, k), l)
        out = eng.ct(batch=a.batch)
        lib.hb_prof_enable(ctx, 1)
        for _ in range(a.reps):
            eng.he_mul(x, y, keys, out=out)
            eng.he_rs(out)
            eng.he_rot(out, 1, keys)
        eng.sync()
        rep = prof_report(eng)
        tot = sum(v["ms"] for v in rep.values())
        print(f"he_mul+he_rs+he_rot logn={a.logn} logq={a.logq} batch={a.batch}: {tot / a.reps:.2f} ms per pass")
        for kname, v in sorted(rep.items(), key=lambda kv: -kv[1]["ms"]):
            print(f"  {kname}: {v['launches']} launches, {v['ms'] / v['launches']:.4f} ms each, share {v['ms'] / tot:.3f}")
    eng.close()


if __name__ == "__main__":
    main()
