"""TEST INFRASTRUCTURE ONLY -- ctypes driver for the compiled, unmodified reference.

Loads ``oracle/_ref/libgpqhe.so`` (built by ``oracle/build_ref.sh`` from the sources under
/root/reference, never copied) and exposes the GPQHE C API (GPQHE/src/gpqhe.h:99-175) plus the
internal symbols the limb-level parity tests need (``ntt``, ``invntt``, ``rns_decompose``,
``poly_rns2mpi``, ``poly_rot``; default visibility).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / reference arm may
import this module.  The product (``hectr_b200``) never does.

The reference keeps ONE process-global context (GPQHE/src/precomp.c:41-47) and a process-global
deterministic RNG (-DSUPERCOP, GPQHE/src/rng.c:32-78); ``Ref`` is therefore a singleton that is
re-initialised when the parameters change.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REFDIR = os.path.join(HERE, "_ref")

MPI = C.c_void_p
u64p = C.POINTER(C.c_uint64)


class RnsCtx(C.Structure):
    pass


RnsCtx._fields_ = [  # GPQHE/src/poly.h:28-41
    ("dim", C.c_uint), ("p", C.c_uint64), ("pinv_mont", C.c_uint64), ("pinv_barr", C.c_uint64),
    ("ninv", C.c_uint64), ("zetas", u64p), ("zetas_inv", u64p), ("P", MPI), ("P_2", MPI),
    ("phat", C.POINTER(MPI)), ("phat_invmp", u64p), ("next", C.POINTER(RnsCtx)),
]


class RingCtx(C.Structure):  # GPQHE/src/poly.h:43-47
    _fields_ = [("cyc_group", C.POINTER(C.c_uint)), ("zetas", C.c_void_p), ("cM", C.c_double)]


class PolyCtx(C.Structure):  # GPQHE/src/poly.h:49-65 (u128 members are 16-byte aligned)
    _fields_ = [
        ("logn", C.c_uint), ("n", C.c_uint), ("m", C.c_uint), ("logqub", C.c_uint), ("logq", C.c_uint),
        ("_pad0", C.c_uint), ("q", MPI), ("logR", C.c_uint), ("_pad1", C.c_uint * 3),
        ("R", C.c_uint64 * 2), ("Rsub1", C.c_uint64 * 2), ("dimub", C.c_uint), ("_pad2", C.c_uint),
        ("rns", C.POINTER(RnsCtx)), ("ring", RingCtx),
    ]


class BndCtx(C.Structure):  # GPQHE/src/gpqhe.h:28-33
    _fields_ = [("Bclean", C.c_double), ("Brs", C.c_double), ("Bks", C.c_double), ("Bmult", C.POINTER(C.c_double))]


class HeCtx(C.Structure):  # GPQHE/src/gpqhe.h:35-50
    _fields_ = [
        ("P", MPI), ("PqL", MPI), ("q", C.POINTER(MPI)), ("qh", C.POINTER(MPI)), ("p", MPI),
        ("slots", C.c_uint), ("Delta", C.c_double), ("L", C.c_uint), ("dim", C.c_uint), ("dimevk", C.c_uint),
        ("bnd", BndCtx), ("Bmult", C.c_double),
    ]


class PolyMpi(C.Structure):  # poly.h:69-72
    _fields_ = [("coeffs", C.POINTER(MPI))]


class PolyRns(C.Structure):  # poly.h:74-77
    _fields_ = [("coeffs", u64p)]


class HeCt(C.Structure):  # gpqhe.h:84-91
    _fields_ = [("l", C.c_uint), ("nu", C.c_double), ("B", C.c_double), ("c0", PolyMpi), ("c1", PolyMpi)]


class HePt(C.Structure):  # gpqhe.h:93-97
    _fields_ = [("nu", C.c_double), ("m", PolyMpi)]


class HeEvk(C.Structure):  # gpqhe.h:78-82
    _fields_ = [("p0", PolyRns), ("p1", PolyRns)]


class HePk(C.Structure):  # gpqhe.h:72-76
    _fields_ = [("p0", PolyMpi), ("p1", PolyMpi)]


def available() -> bool:
    return os.path.exists(os.path.join(REFDIR, "libgpqhe.so")) and os.path.exists(
        os.path.join(REFDIR, "libref_helpers.so"))


_LIBS = {}


def _load(opt: str = "Og"):
    if opt in _LIBS:
        return _LIBS[opt]
    name = "libgpqhe.so" if opt == "Og" else "libgpqhe_O2.so"
    gcry = C.CDLL("libgcrypt.so.20", mode=C.RTLD_GLOBAL)
    lib = C.CDLL(os.path.join(REFDIR, name), mode=C.RTLD_GLOBAL)
    hlp = C.CDLL(os.path.join(REFDIR, "libref_helpers.so"), mode=C.RTLD_GLOBAL)
    hlp.ref_pow2.restype = MPI
    hlp.ref_mpi_to_words.restype = C.c_int
    gcry.gcry_mpi_release.argtypes = [MPI]
    gcry.gcry_mpi_new.restype = MPI
    gcry.gcry_mpi_get_nbits.argtypes = [MPI]
    _LIBS[opt] = (lib, hlp, gcry)
    return _LIBS[opt]


class Ref:
    """One live reference context.  ``Ref.get(logn, logq, slots, logDelta)`` (re)initialises."""

    _cur = None

    @classmethod
    def get(cls, logn, logq, slots, logDelta, opt="Og"):
        key = (logn, logq, slots, logDelta, opt)
        if cls._cur is not None and cls._cur.key == key:
            return cls._cur
        if cls._cur is not None:
            cls._cur.close()
        cls._cur = cls(*key)
        return cls._cur

    def __init__(self, logn, logq, slots, logDelta, opt="Og"):
        self.key = (logn, logq, slots, logDelta, opt)
        self.lib, self.hlp, self.gcry = _load(opt)
        lib = self.lib
        q = MPI(self.hlp.ref_pow2(C.c_uint(logq)))
        lib.hectx_init(C.c_uint(logn), q, C.c_uint(slots), C.c_uint64(1 << logDelta))
        self.gcry.gcry_mpi_release(q)
        self.polyctx = PolyCtx.in_dll(lib, "polyctx")
        self.hectx = HeCtx.in_dll(lib, "hectx")
        self.logn, self.n, self.logq, self.slots, self.logDelta = logn, 1 << logn, logq, slots, logDelta
        self.L, self.dim, self.dimevk, self.dimub = (self.hectx.L, self.hectx.dim, self.hectx.dimevk,
                                                      self.polyctx.dimub)
        self._nodes = []
        node = self.polyctx.rns
        while node:
            self._nodes.append(node.contents)
            node = node.contents.next
        lib.poly_rot.argtypes = [C.POINTER(PolyMpi), C.POINTER(PolyMpi), C.c_size_t]

    def close(self):
        self.lib.hectx_exit()
        Ref._cur = None

    # ---- tables -------------------------------------------------------------------------------
    def node(self, d):
        return self._nodes[d]

    def primes(self):
        return [int(nd.p) for nd in self._nodes]

    def zetas(self, d, inv=False):
        nd = self._nodes[d]
        src = nd.zetas_inv if inv else nd.zetas
        return np.ctypeslib.as_array(src, shape=(self.n,)).copy()

    def mpi_int(self, m) -> int:
        """Python int of a libgcrypt MPI (any size up to 4096 bits)."""
        W = (self.gcry.gcry_mpi_get_nbits(MPI(m) if not isinstance(m, MPI) else m) + 64) // 64 + 1
        arr = (MPI * 1)(m)
        out = np.zeros(W, dtype=np.uint64)
        rc = self.hlp.ref_mpi_to_words(arr, C.c_uint(1), C.c_uint(W), out.ctypes.data_as(u64p))
        assert rc == 0
        v = 0
        for w in range(W):
            v |= int(out[w]) << (64 * w)
        if v >> (64 * W - 1):
            v -= 1 << (64 * W)
        return v

    def q_level(self, l) -> int:
        return self.mpi_int(self.hectx.q[l])

    # ---- containers ---------------------------------------------------------------------------
    def words_per_level(self, l) -> int:
        return (self.q_level(l).bit_length() + 63) // 64

    def alloc_ct(self):
        ct = HeCt()
        self.lib.he_alloc_ct(C.byref(ct))
        return ct

    def free_ct(self, ct):
        self.lib.he_free_ct(C.byref(ct))

    def alloc_pt(self):
        pt = HePt()
        self.lib.he_alloc_pt(C.byref(pt))
        return pt

    def alloc_poly(self):
        p = PolyMpi()
        self.lib.poly_mpi_alloc(C.byref(p))
        return p

    def poly_to_words(self, poly: PolyMpi, W: int) -> np.ndarray:
        out = np.zeros((W, self.n), dtype=np.uint64)
        rc = self.hlp.ref_mpi_to_words(poly.coeffs, C.c_uint(self.n), C.c_uint(W), out.ctypes.data_as(u64p))
        if rc:
            raise ValueError(f"coefficient {rc - 1} does not fit {W} words")
        return out

    def words_to_poly(self, poly: PolyMpi, words: np.ndarray):
        words = np.ascontiguousarray(words, dtype=np.uint64)
        W = words.shape[0]
        self.hlp.ref_words_to_mpi(poly.coeffs, C.c_uint(self.n), C.c_uint(W), words.ctypes.data_as(u64p))

    def ct_to_words(self, ct: HeCt, W=None) -> np.ndarray:
        """-> uint64[2][W][N], planar little-endian two's complement."""
        W = W or self.words_per_level(ct.l)
        return np.stack([self.poly_to_words(ct.c0, W), self.poly_to_words(ct.c1, W)])

    def ct_from_words(self, words: np.ndarray, l: int, nu=1.0, B=1.0) -> HeCt:
        ct = self.alloc_ct()
        ct.l, ct.nu, ct.B = l, nu, B
        self.words_to_poly(ct.c0, words[0])
        self.words_to_poly(ct.c1, words[1])
        return ct

    def copy_ct(self, src: HeCt) -> HeCt:
        dst = self.alloc_ct()
        self.lib.he_copy_ct(C.byref(dst), C.byref(src))
        return dst

    # ---- keys / client side ---------------------------------------------------------------------
    def keypair(self):
        sk = PolyMpi()
        pk = HePk()
        self.lib.he_alloc_sk(C.byref(sk))
        self.lib.he_alloc_pk(C.byref(pk))
        self._quiet(lambda: self.lib.he_keypair(C.byref(pk), C.byref(sk)))
        return pk, sk

    def _quiet(self, fn):
        """The reference printf()s progress from keygen (he-kem.c:49,70,127,145,160)."""
        import sys
        sys.stdout.flush()
        fd = os.dup(1)
        nul = os.open(os.devnull, os.O_WRONLY)
        os.dup2(nul, 1)
        try:
            fn()
            C.CDLL(None).fflush(None)
        finally:
            os.dup2(fd, 1)
            os.close(nul)
            os.close(fd)

    def genrk(self, sk):
        rk = (HeEvk * self.slots)()
        for i in range(self.slots):
            self.lib.he_alloc_evk(C.byref(rk[i]))
        self._quiet(lambda: self.lib.he_genrk(rk, C.byref(sk)))
        return rk

    def genrlk(self, sk):
        rlk = HeEvk()
        self.lib.he_alloc_evk(C.byref(rlk))
        self._quiet(lambda: self.lib.he_genrlk(C.byref(rlk), C.byref(sk)))
        return rlk

    def genck(self, sk):
        ck = HeEvk()
        self.lib.he_alloc_evk(C.byref(ck))
        self._quiet(lambda: self.lib.he_genck(C.byref(ck), C.byref(sk)))
        return ck

    def evk_to_array(self, evk: HeEvk) -> np.ndarray:
        """-> uint64[2][dimevk][N] (NTT/RNS form exactly as he_genswk leaves it, he-kem.c:103-110)."""
        shape = (self.dimevk, self.n)
        return np.stack([np.ctypeslib.as_array(evk.p0.coeffs, shape=shape).copy(),
                         np.ctypeslib.as_array(evk.p1.coeffs, shape=shape).copy()])

    def evk_from_array(self, arr: np.ndarray) -> HeEvk:
        """uint64[2][dimevk][N] -> a reference he_evk_t holding those limbs (keys made elsewhere, e.g. by hb_he_genswk)."""
        arr = np.ascontiguousarray(arr, dtype=np.uint64)
        assert arr.shape == (2, self.dimevk, self.n), arr.shape
        evk = HeEvk()
        self.lib.he_alloc_evk(C.byref(evk))
        C.memmove(evk.p0.coeffs, arr[0].ctypes.data, arr[0].nbytes)
        C.memmove(evk.p1.coeffs, arr[1].ctypes.data, arr[1].nbytes)
        return evk

    def rk_from_array(self, arr: np.ndarray):
        """uint64[slots][2][dimevk][N] -> he_evk_t[slots]"""
        rk = (HeEvk * self.slots)()
        for i in range(self.slots):
            self.lib.he_alloc_evk(C.byref(rk[i]))
            a = np.ascontiguousarray(arr[i], dtype=np.uint64)
            C.memmove(rk[i].p0.coeffs, a[0].ctypes.data, a[0].nbytes)
            C.memmove(rk[i].p1.coeffs, a[1].ctypes.data, a[1].nbytes)
        return rk

    def encode(self, m) -> HePt:
        m = np.ascontiguousarray(m, dtype=np.complex128)
        assert m.shape == (self.slots,)
        pt = self.alloc_pt()
        self.lib.he_ecd(C.byref(pt), m.ctypes.data_as(C.c_void_p))
        return pt

    def decode(self, pt: HePt) -> np.ndarray:
        m = np.zeros(self.slots, dtype=np.complex128)
        self.lib.he_dcd(m.ctypes.data_as(C.c_void_p), C.byref(pt))
        return m

    def enc_pk(self, pt: HePt, pk: HePk) -> HeCt:
        ct = self.alloc_ct()
        self.lib.he_enc_pk(C.byref(ct), C.byref(pt), C.byref(pk))
        return ct

    def dec(self, ct: HeCt, sk: PolyMpi) -> HePt:
        pt = self.alloc_pt()
        self.lib.he_dec(C.byref(pt), C.byref(ct), C.byref(sk))
        return pt

    # ---- samplers and the bare polynomial product (client-side parity: he-encrypt.c) -------------------------
    def _sample(self, fn) -> np.ndarray:
        p = self.alloc_poly()
        fn(C.byref(p))
        out = self.poly_to_words(p, 1)[0].view(np.int64).copy()
        self.lib.poly_mpi_free(C.byref(p))
        return out

    def sample_zo(self) -> np.ndarray:
        """sample_zo (sample.c:123-131) -> int64[N]; advances the deterministic stream like he_enc_pk's first draw."""
        return self._sample(self.lib.sample_zo)

    def sample_error(self) -> np.ndarray:
        """sample_error (sample.c:74-82) -> int64[N]."""
        return self._sample(self.lib.sample_error)

    def sample_uniform_PqL(self) -> PolyMpi:
        """sample_uniform(., P q_L) (sample.c:133-141), he_genswk's draw of swk.p1 (he-kem.c:94)."""
        p = self.alloc_poly()
        self.lib.sample_uniform(C.byref(p), self.hectx.PqL)
        return p

    def poly_mul_PqL(self, a: PolyMpi, b: PolyMpi, dim: int) -> list:
        """poly_mul(a, b, dim, P q_L) (poly.c:84-107) -> python ints (centred mod P q_L)."""
        r = self.alloc_poly()
        self.lib.poly_mul(C.byref(r), C.byref(a), C.byref(b), C.c_uint(dim), self.hectx.PqL)
        out = [self.mpi_int(r.coeffs[i]) for i in range(self.n)]
        self.lib.poly_mpi_free(C.byref(r))
        return out

    def poly_mul_words(self, a: PolyMpi, b_words: np.ndarray, dim: int, logm: int, W: int) -> np.ndarray:
        """poly_mul(a, b, dim, 2^logm) (poly.c:84-107) -> words[W][N]."""
        b = self.alloc_poly()
        self.words_to_poly(b, b_words)
        r = self.alloc_poly()
        q = MPI(self.hlp.ref_pow2(C.c_uint(logm)))
        self.lib.poly_mul(C.byref(r), C.byref(a), C.byref(b), C.c_uint(dim), q)
        self.gcry.gcry_mpi_release(q)
        out = self.poly_to_words(r, W)
        self.lib.poly_mpi_free(C.byref(b))
        self.lib.poly_mpi_free(C.byref(r))
        return out

    # ---- evaluation (each returns / mutates reference structs) --------------------------------------
    def add(self, a, b):
        r = self.alloc_ct()
        self.lib.he_add(C.byref(r), C.byref(a), C.byref(b))
        return r

    def sub(self, a, b):
        r = self.alloc_ct()
        self.lib.he_sub(C.byref(r), C.byref(a), C.byref(b))
        return r

    def neg(self, a):
        self.lib.he_neg(C.byref(a))
        return a

    def moddown(self, a):
        self.lib.he_moddown(C.byref(a))
        return a

    def rs(self, a):
        self.lib.he_rs(C.byref(a))
        return a

    def mulpt(self, a, pt):
        r = self.alloc_ct()
        self.lib.he_mulpt(C.byref(r), C.byref(a), C.byref(pt))
        return r

    def mul(self, a, b, rlk):
        r = self.alloc_ct()
        self.lib.he_mul(C.byref(r), C.byref(a), C.byref(b), C.byref(rlk))
        return r

    def rot(self, a, r, rk):
        self.lib.he_rot(C.byref(a), C.c_int(r), rk)
        return a

    def conj(self, a, ck):
        self.lib.he_conj(C.byref(a), C.byref(ck))
        return a

    def gemv(self, A, a, rk):
        A = np.ascontiguousarray(A, dtype=np.complex128)
        assert A.shape == (self.slots, self.slots)
        r = self.alloc_ct()
        self.lib.he_gemv(C.byref(r), A.ctypes.data_as(C.c_void_p), C.byref(a), rk)
        return r

    # ---- he-add.c:80-123, he-encode.c:119-125, he-algo.c:95-548 -----------------
    def const_pt(self, num: complex, pt=None):
        """he_const_pt takes a _Complex double by value: two SSE eightbytes, i.e. two double arguments (SysV x86-64)."""
        pt = pt or self.alloc_pt()
        num = complex(num)
        self.lib.he_const_pt(C.byref(pt), C.c_double(num.real), C.c_double(num.imag))
        return pt

    def addpt(self, a, pt):
        r = self.alloc_ct()
        self.lib.he_addpt(C.byref(r), C.byref(a), C.byref(pt))
        return r

    def subpt(self, a, pt):
        r = self.alloc_ct()
        self.lib.he_subpt(C.byref(r), C.byref(a), C.byref(pt))
        return r

    def _algo(self, name, *args):
        r = self.alloc_ct()
        self._quiet(lambda: getattr(self.lib, name)(C.byref(r), *args))
        return r

    def he_sum(self, a, rk):
        return self._algo("he_sum", C.byref(a), rk)

    def he_idx(self, a, idx, rk):
        return self._algo("he_idx", C.byref(a), C.c_uint(idx), rk)

    def he_nrm2(self, a, rlk, ck, rk):
        return self._algo("he_nrm2", C.byref(a), C.byref(rlk), C.byref(ck), rk)

    def he_inv(self, a, rlk, iters):
        return self._algo("he_inv", C.byref(a), C.byref(rlk), C.c_uint(iters))

    def he_sqrt(self, a, rlk, iters):
        return self._algo("he_sqrt", C.byref(a), C.byref(rlk), C.c_uint(iters))

    def he_sigmoid(self, a, rlk):
        return self._algo("he_sigmoid", C.byref(a), C.byref(rlk))

    def he_log(self, a, rlk):
        return self._algo("he_log", C.byref(a), C.byref(rlk))

    def he_exp(self, coef: complex, a, rlk, iters):
        coef = complex(coef)
        return self._algo("he_exp", C.c_double(coef.real), C.c_double(coef.imag), C.byref(a), C.byref(rlk), C.c_uint(iters))

    def he_cmp(self, a, b, rlk, iters, alpha):
        return self._algo("he_cmp", C.byref(a), C.byref(b), C.byref(rlk), C.c_uint(iters), C.c_uint(alpha))

    def he_cmppt(self, a, pt, rlk, iters, alpha):
        return self._algo("he_cmppt", C.byref(a), C.byref(pt), C.byref(rlk), C.c_uint(iters), C.c_uint(alpha))

    # ---- limb-level internals ---------------------------------------------------------------------
    def ntt(self, a: np.ndarray, d: int, inverse=False) -> np.ndarray:
        a = np.ascontiguousarray(a, dtype=np.uint64).copy()
        fn = self.lib.invntt if inverse else self.lib.ntt
        fn(a.ctypes.data_as(u64p), C.byref(self._nodes[d]))
        return a

    def rns_decompose(self, poly: PolyMpi, d: int) -> np.ndarray:
        out = np.zeros(self.n, dtype=np.uint64)
        self.lib.rns_decompose(out.ctypes.data_as(u64p), poly.coeffs, C.byref(self._nodes[d]))
        return out

    def rns2mpi(self, limbs: np.ndarray, logm: int, W: int) -> np.ndarray:
        """poly_rns2mpi(limbs[D][N]) with the D-limb node and modulus 2^logm -> words[W][N]."""
        limbs = np.ascontiguousarray(limbs, dtype=np.uint64)
        D = limbs.shape[0]
        r = self.alloc_poly()
        rhat = PolyRns(limbs.ctypes.data_as(u64p))
        q = MPI(self.hlp.ref_pow2(C.c_uint(logm)))
        self.lib.poly_rns2mpi(C.byref(r), C.byref(rhat), C.byref(self._nodes[D - 1]), q)
        self.gcry.gcry_mpi_release(q)
        out = self.poly_to_words(r, W)
        self.lib.poly_mpi_free(C.byref(r))
        return out

    def poly_rot(self, words: np.ndarray, rot: int) -> np.ndarray:
        a, r = self.alloc_poly(), self.alloc_poly()
        self.words_to_poly(a, words)
        self.lib.poly_rot(C.byref(r), C.byref(a), C.c_size_t(rot))
        out = self.poly_to_words(r, words.shape[0])
        self.lib.poly_mpi_free(C.byref(a))
        self.lib.poly_mpi_free(C.byref(r))
        return out
