"""TEST INFRASTRUCTURE ONLY -- an independent big-integer restatement of the reference's CKKS evaluation path.

Python integers, no numpy, no ctypes: nothing here shares code with ``hectr_b200`` or with ``oracle/ref.py``
(the compiled, unmodified reference).  Purpose: a second, readable statement of the algorithm that
  * is pinned (tests/test_oracle_restate.py) against the reference's own known answers -- the two logn=7 primes
    and the exact product of GPQHE/tests/polymul.gp:1-23 -- and against tests/golden/*.npz, vectors produced by
    the compiled reference (tests/golden/make_golden.py);
  * travels to the GPU box as source, so the parity tests still have a checker where /root/reference is absent.
It is slow (pure-Python loops): use it at N <= 2^10 or on a few coefficients.  Only ``tests/`` may import it.

Each function cites the reference lines it restates (paths relative to /root/reference/GPQHE/src).
A polynomial is a list of N Python ints; an RNS limb is a list of N ints in [0, p]; a ciphertext is
``{"l": level, "c0": poly, "c1": poly}``; an evaluation key is ``(p0, p1)``, each a list of dimevk limbs (NTT form).
"""
from __future__ import annotations

LOGP = 59          # params.h:26-28  GPQHE_LOGP
ROT = 5            # params.h GPQHE_ROT
R64 = 1 << 64      # Montgomery radix, precomp.c:362-364
STD_LOGQ = {10: 27, 11: 54, 12: 109, 13: 218, 14: 438, 15: 881}   # he_std_params, precomp.c:53-64; other logn: no cap


# ------------------------------------------------------------------------------------------------------
# scalar helpers
# ------------------------------------------------------------------------------------------------------
def is_prime(p: int) -> bool:
    """Deterministic Miller-Rabin for 64-bit inputs (the reference uses 50 rand() witnesses, precomp.c:153-191;
    for 64-bit integers the first twelve primes as witnesses decide the same predicate exactly)."""
    if p < 2:
        raise ValueError("neither prime nor composite")
    small = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)
    for s in small:
        if p % s == 0:
            return p == s
    u, t = p - 1, 0
    while u % 2 == 0:
        u //= 2
        t += 1
    for a in small:
        x = pow(a, u, p)
        if x in (1, p - 1):
            continue
        for _ in range(t - 1):
            x = x * x % p
            if x == p - 1:
                break
        else:
            return False
    return True


def factor_distinct(n: int):
    """Distinct prime factors by trial division (precomp.c:194-204 keeps multiplicities; only the set matters)."""
    out, f = [], 2
    while f * f <= n:
        if n % f == 0:
            out.append(f)
            while n % f == 0:
                n //= f
        f += 1 if f == 2 else 2
    if n > 1:
        out.append(n)
    return out


def generator(p: int) -> int:
    """Least generator of (Z/p)^* (precomp.c:206-226)."""
    phi = p - 1
    fs = factor_distinct(phi)
    g = 2
    while any(pow(g, phi // f, p) == 1 for f in fs):
        g += 1
    return g


def bitrev(i: int, bits: int) -> int:
    r = 0
    for _ in range(bits):
        r = (r << 1) | (i & 1)
        i >>= 1
    return r


def smod(r: int, q: int) -> int:
    """types.c:108-113: r mod q, then r >= floor(q/2) -> r - q.  Range [-ceil(q/2), floor(q/2))."""
    r %= q
    return r - q if r >= q // 2 else r


def rdiv(a: int, m: int) -> int:
    """types.c:115-128: floor(a/m) + [a mod m > floor(m/2)]  (floor division; ties go down)."""
    assert m > 0
    q, r = divmod(a, m)
    return q + 1 if r > m // 2 else q


def montgomery_reduce(a: int, p: int, pinv: int) -> int:
    """reduce.c:59-66 (a < 2^128): hi - floor(((a mod R) pinv mod R) p / R), plus p when negative."""
    lo, hi = a % R64, a >> 64
    u = lo * pinv % R64
    t = (u * p) >> 64
    return hi - t + p if hi < t else hi - t


def barrett_reduce(a: int, p: int, pinv_barr: int) -> int:
    """reduce.c:88-106: one conditional subtraction after the quotient estimate with shift 2*nbits(p)-64."""
    lo, hi = a % R64, a >> 64
    t = ((lo * pinv_barr) >> 64) + hi * pinv_barr
    t >>= 2 * p.bit_length() - 64
    r = (a - t * p) % R64
    return r if r < p else r - p


# ------------------------------------------------------------------------------------------------------
# context: prime chain, twiddles, CRT constants, level table   (precomp.c:244-293, 328-409, 431-453)
# ------------------------------------------------------------------------------------------------------
class Prime:
    def __init__(self, p: int, logn: int):
        n = 1 << logn
        self.p = p
        self.pinv_mont = pow(p, -1, R64)                       # reduce.c:40-52 montgomery_inv
        self.pinv_barr = (1 << (2 * p.bit_length())) // p      # reduce.c:74-77
        self.ninv = pow(n, p - 2, p) * R64 % p                 # precomp.c:248
        root = pow(generator(p), (p - 1) // (2 * n), p)        # precomp.c:228-242
        rinv = pow(root, p - 2, p)
        self.root = root
        self.zetas = [0] * n
        self.zetas_inv = [0] * n
        pw = pwi = 1
        for i in range(n):                                     # precomp.c:255-263
            j = bitrev(i, logn)
            self.zetas[j] = pw * R64 % p
            self.zetas_inv[j] = pwi * R64 % p
            pw = pw * root % p
            pwi = pwi * rinv % p


class Ctx:
    """hectx_init(logn, q = 2^logq, slots, Delta = 2^logDelta)."""

    def __init__(self, logn: int, logq: int, slots: int, logDelta: int, dimub: int | None = None):
        self.logn, self.n, self.logq, self.slots, self.logDelta = logn, 1 << logn, logq, slots, logDelta
        q = 1 << logq
        self.Delta = 1 << logDelta
        logqub = STD_LOGQ.get(logn, logq)                      # precomp.c:53-64 (128-bit classical table), :336-340
        if logq > logqub:
            raise ValueError("The input modulus q is too large. Must guarantee log(q)<=log(qub).")   # precomp.c:344-350
        full = (1 + logn + 4 * logqub) // LOGP + 1             # precomp.c:357
        self.dimub = full if dimub is None else min(dimub, full)
        self.primes = []
        p = (1 << LOGP) + 1                                    # precomp.c:358
        while len(self.primes) < self.dimub:                   # precomp.c:372-376
            p += 2 * self.n
            if is_prime(p):
                self.primes.append(Prime(p, logn))
        # "use the first D limbs": P_D, floor(P_D/2), P_D/p_d, (P_D/p_d)^-1 mod p_d     (precomp.c:266-293)
        self.P, self.phat, self.phat_inv = [1], [[]], [[]]
        for D in range(1, self.dimub + 1):
            PD = self.P[-1] * self.primes[D - 1].p
            self.P.append(PD)
            self.phat.append([PD // self.primes[d].p for d in range(D)])
            self.phat_inv.append([pow(PD // self.primes[d].p % self.primes[d].p, self.primes[d].p - 2, self.primes[d].p)
                                  for d in range(D)])
        # level table (precomp.c:386-409); logq / logDelta is an integer division
        self.L = logq // logDelta
        self.q = [0] * (self.L + 1)
        x = q
        for l in range(self.L, -1, -1):
            self.q[l] = x
            x //= self.Delta
        self.dim = (self.q[self.L].bit_length() + logn) // LOGP + 1
        if self.dim <= self.dimub:
            self.Pks = self.P[self.dim]
            self.PqL = self.Pks * q
            self.dimevk = (self.q[self.L].bit_length() + self.PqL.bit_length() + logn) // LOGP + 1

    # limb counts (he-automorphism.c:52, he-mult.c:51,99,168)
    def dim_keyswitch(self, l):
        return (self.q[l].bit_length() + self.PqL.bit_length() + self.logn) // LOGP + 1

    def dim_mul(self, l):
        return (self.q[l].bit_length() * 2 + self.logn) // LOGP + 1

    def dim_mulpt(self, l, log2nu: float):
        return int((self.q[l].bit_length() + log2nu + self.logn) / LOGP + 1)

    # ---- transforms (ntt.c:37-73) ------------------------------------------------------------------------
    def ntt(self, a, d):
        pr = self.primes[d]
        p, pinv, z = pr.p, pr.pinv_mont, pr.zetas
        a = list(a)
        n, k, length = self.n, 1, self.n // 2
        while length >= 1:
            for start in range(0, n, 2 * length):
                zeta = z[k]
                k += 1
                for j in range(start, start + length):
                    t = montgomery_reduce(a[j + length] * zeta, p, pinv)
                    x = a[j]
                    a[j + length] = x - t if x >= t else x - t + p
                    a[j] = x + t if x <= p - t else x + t - p      # a sum of exactly p stays p
            length >>= 1
        return a

    def invntt(self, a, d):
        pr = self.primes[d]
        p, pinv, z = pr.p, pr.pinv_mont, pr.zetas_inv
        a = list(a)
        n, length = self.n, 1
        while length <= n // 2:
            k = n // (2 * length)
            for start in range(0, n, 2 * length):
                zeta = z[k]
                k += 1
                for j in range(start, start + length):
                    t, y = a[j], a[j + length]
                    a[j] = t + y if y <= p - t else t + y - p
                    dlt = t - y if y <= t else t - y + p
                    a[j + length] = montgomery_reduce(dlt * zeta, p, pinv)
            length <<= 1
        return [montgomery_reduce(x * pr.ninv, p, pinv) for x in a]

    # ---- RNS <-> integers (rns.c:37-75, poly.c:109-120) -------------------------------------------------------
    def decompose(self, poly, d):
        p = self.primes[d].p
        return [x % p for x in poly]                           # gcry_mpi_mod: non-negative

    def reconstruct(self, limbs, m):
        """poly_rns2mpi: CRT over len(limbs) primes, centred mod P_D, then centred mod m."""
        D = len(limbs)
        PD = self.P[D]
        cs = [self.phat[D][d] * self.phat_inv[D][d] % PD for d in range(D)]
        out = []
        for i in range(self.n):
            x = 0
            for d in range(D):
                x = (x + limbs[d][i] * cs[d]) % PD
            out.append(smod(smod(x, PD), m))
        return out

    def rns_mul(self, a, b, d):
        pr = self.primes[d]
        return [barrett_reduce(x * y, pr.p, pr.pinv_barr) for x, y in zip(a, b)]

    def rns_add(self, a, b, d):
        pr = self.primes[d]
        return [barrett_reduce(x + y, pr.p, pr.pinv_barr) for x, y in zip(a, b)]

    def poly_mul(self, a, b, D, m):
        """poly.c:84-107."""
        limbs = []
        for d in range(D):
            limbs.append(self.invntt(self.rns_mul(self.ntt(self.decompose(a, d), d), self.ntt(self.decompose(b, d), d), d), d))
        return self.reconstruct(limbs, m)

    # ---- automorphisms (poly.c:263-283) ----------------------------------------------------------------------
    def poly_rot(self, a, rot):
        n, m = self.n, 2 * self.n
        power = pow(ROT, rot, 1 << 64)                         # wrapping size_t; 2N divides 2^64
        r = [0] * n
        for i in range(n):
            k = i * power % m
            if k < n:
                r[k] = a[i]
            else:
                r[k - n] = -a[i]
        return r

    def poly_conj(self, a):
        n = self.n
        return [a[0]] + [-a[n - i] for i in range(1, n)]

    # ---- ciphertext arithmetic ---------------------------------------------------------------------------------
    def he_add(self, a, b, sign=+1):
        """he-add.c:32-76: addm/subm then smod."""
        assert a["l"] == b["l"]
        q = self.q[a["l"]]
        return {"l": a["l"], "c0": [smod((x + sign * y) % q, q) for x, y in zip(a["c0"], b["c0"])],
                "c1": [smod((x + sign * y) % q, q) for x, y in zip(a["c1"], b["c1"])]}

    def he_sub(self, a, b):
        return self.he_add(a, b, -1)

    def he_neg(self, a):
        """he-add.c:126-142."""
        q = self.q[a["l"]]
        return {"l": a["l"], "c0": [smod(-x, q) for x in a["c0"]], "c1": [smod(-x, q) for x in a["c1"]]}

    def he_rs(self, a):
        """he-rescale.c:33-54."""
        l = a["l"] - 1
        q = self.q[l]
        return {"l": l, "c0": [smod(rdiv(x, self.Delta), q) for x in a["c0"]],
                "c1": [smod(rdiv(x, self.Delta), q) for x in a["c1"]]}

    def he_moddown(self, a):
        """he-rescale.c:56-70."""
        l = a["l"] - 1
        q = self.q[l]
        return {"l": l, "c0": [smod(x, q) for x in a["c0"]], "c1": [smod(x, q) for x in a["c1"]]}

    def he_mulpt(self, a, pt, log2nu: float):
        """he-mult.c:159-196."""
        l = a["l"]
        D = self.dim_mulpt(l, log2nu)
        out = {"l": l}
        pthat = [self.ntt(self.decompose(pt, d), d) for d in range(D)]
        for nm in ("c0", "c1"):
            limbs = [self.invntt(self.rns_mul(self.ntt(self.decompose(a[nm], d), d), pthat[d], d), d) for d in range(D)]
            out[nm] = self.reconstruct(limbs, self.q[l])
        return out

    def _switch(self, l, d_switched, evk):
        """The common core of he_swk / he_relin: both key products, reconstruct mod P q_l, rounded division by P."""
        ql = self.q[l]
        D = self.dim_keyswitch(l)
        outs = []
        hats = [self.ntt(self.decompose(d_switched, d), d) for d in range(D)]
        for h in range(2):
            limbs = [self.invntt(self.rns_mul(hats[d], evk[h][d], d), d) for d in range(D)]
            x = self.reconstruct(limbs, self.Pks * ql)
            outs.append([rdiv(v, self.Pks) for v in x])
        return outs

    def he_swk(self, l, d0, d1, evk):
        """he-automorphism.c:40-85: c0 = smod(rdiv(d1 k0, P) + d0, q_l); c1 = smod(rdiv(d1 k1, P), q_l)."""
        ql = self.q[l]
        c0, c1 = self._switch(l, d1, evk)
        return {"l": l, "c0": [smod((x + y) % ql, ql) for x, y in zip(c0, d0)], "c1": [smod(x, ql) for x in c1]}

    def he_rot(self, a, rot, rk):
        """he-automorphism.c:103-115 (rot = 0 still key-switches with rk[0])."""
        return self.he_swk(a["l"], self.poly_rot(a["c0"], rot), self.poly_rot(a["c1"], rot), rk[rot])

    def he_conj(self, a, ck):
        """he-automorphism.c:87-101."""
        return self.he_swk(a["l"], self.poly_conj(a["c0"]), self.poly_conj(a["c1"]), ck)

    def he_mul(self, a, b, rlk):
        """he-mult.c:88-156 with he_relin :40-85 (relin adds d0 AND d1)."""
        assert a["l"] == b["l"]
        l = a["l"]
        ql = self.q[l]
        D = self.dim_mul(l)
        d0h, d1h, d2h = [], [], []
        for d in range(D):
            a0, a1 = self.ntt(self.decompose(a["c0"], d), d), self.ntt(self.decompose(a["c1"], d), d)
            b0, b1 = self.ntt(self.decompose(b["c0"], d), d), self.ntt(self.decompose(b["c1"], d), d)
            d0h.append(self.invntt(self.rns_mul(a0, b0, d), d))
            d2h.append(self.invntt(self.rns_mul(a1, b1, d), d))
            d1h.append(self.rns_add(self.invntt(self.rns_mul(a0, b1, d), d), self.invntt(self.rns_mul(a1, b0, d), d), d))
        d0, d1, d2 = self.reconstruct(d0h, ql), self.reconstruct(d1h, ql), self.reconstruct(d2h, ql)
        c0, c1 = self._switch(l, d2, rlk)
        return {"l": l, "c0": [smod((x + y) % ql, ql) for x, y in zip(c0, d0)],
                "c1": [smod((x + y) % ql, ql) for x, y in zip(c1, d1)]}

    def he_gemv(self, a, diag_pts, rk):
        """he-algo.c:47-93 given the slots encoded, pre-rotated diagonals diag_pts[i*n1 + j] (he_ecd(zrotdiag(A, i n1 + j,
        -i n1)), nu = Delta).  Baby/giant split exactly as lines 51-55."""
        import math
        S = self.slots
        n1 = int(math.sqrt(S))
        if S != n1 * n1:
            n1 = int(math.sqrt(2 * S))
        n2 = S // n1
        outer = None
        for i in range(n2):
            inner = None
            for j in range(n1):
                r = self.he_mulpt(self.he_rot(a, j, rk), diag_pts[i * n1 + j], float(self.logDelta))
                inner = r if inner is None else self.he_add(inner, r)
            inner = self.he_rot(inner, i * n1, rk)
            outer = inner if outer is None else self.he_add(outer, inner)
        return self.he_rs(outer)

    def ctr_hempc(self, xhat, xr, uhat, ur, Ka_pts, Kb_pts, rk):
        """The ciphertext half of ctr_hempc, /root/reference/src/hempc.c:253-266."""
        du = self.he_neg(self.he_add(self.he_gemv(self.he_sub(xhat, xr), Ka_pts, rk),
                                     self.he_gemv(self.he_sub(uhat, ur), Kb_pts, rk)))
        return self.he_add(self.he_moddown(uhat), du)


def negacyclic_schoolbook(a, b):
    """a*b mod (X^n + 1) over the integers -- what GPQHE/tests/polymul.gp computes with PARI (Mod(a*b, x^128+1))."""
    n = len(a)
    r = [0] * n
    for i, x in enumerate(a):
        if x == 0:
            continue
        for j, y in enumerate(b):
            k = i + j
            if k < n:
                r[k] += x * y
            else:
                r[k - n] -= x * y
    return r
