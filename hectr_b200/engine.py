"""Host-side mirror of the GPQHE evaluation API over the C ABI (``include/hectr_b200.h``).

Method names follow the reference (``he_add``, ``he_rot``, ``he_gemv`` ... GPQHE/src/gpqhe.h:99-175);
arguments are numpy arrays in the layouts the header documents.  Everything here is plumbing:
the arithmetic happens in ``libhectr_b200.so`` on the GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import HbCt, HbEvk, HbInfo, HbPt, HbSk, load_library


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Ciphertext:
    """A batch of ciphertexts resident on the GPU (``hb_ct``; he_ct_t gpqhe.h:84-91)."""

    def __init__(self, eng: "Engine", batch: int):
        self.eng = eng
        self.h = HbCt()
        eng.lib.hb_ct_alloc(eng.ctx, C.byref(self.h), batch)

    l = property(lambda s: s.h.l)
    nu = property(lambda s: s.h.nu)
    B = property(lambda s: s.h.B)
    batch = property(lambda s: s.h.batch)

    def free(self):
        if self.h.words:
            self.eng.lib.hb_ct_free(self.eng.ctx, C.byref(self.h))

    def upload(self, words: np.ndarray, l: int, nu: float = 1.0, B: float = 1.0):
        W = self.eng.words(l)
        words = np.ascontiguousarray(words, dtype=np.uint64)
        assert words.shape == (self.batch, 2, W, self.eng.n), (words.shape, (self.batch, 2, W, self.eng.n))
        self.eng.lib.hb_ct_upload(self.eng.ctx, C.byref(self.h), _ptr(words), l, nu, B)
        self.eng.lib.hb_sync(self.eng.ctx)
        return self

    def download(self) -> np.ndarray:
        out = np.empty((self.batch, 2, self.eng.words(self.l), self.eng.n), dtype=np.uint64)
        self.eng.lib.hb_ct_download(self.eng.ctx, C.byref(self.h), _ptr(out))
        return out


class EvalKeys:
    """An array of evaluation keys on the GPU (rk[slots], or a single rlk / ck)."""

    def __init__(self, eng: "Engine", keys: np.ndarray):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        if keys.ndim == 3:
            keys = keys[None]
        assert keys.shape[1:] == (2, eng.info.dimevk, eng.n), keys.shape
        self.eng = eng
        self.arr = (HbEvk * keys.shape[0])()
        for i in range(keys.shape[0]):
            eng.lib.hb_evk_upload(eng.ctx, C.byref(self.arr[i]), _ptr(keys[i, 0]), _ptr(keys[i, 1]))

    def free(self):
        for e in self.arr:
            if e.p:
                self.eng.lib.hb_evk_free(self.eng.ctx, C.byref(e))


class Engine:
    """One engine context = one ``hectx_init`` (GPQHE/src/precomp.c:431-453) on one GPU."""

    def __init__(self, logn: int, logq: int, slots: int, logDelta: int, device: int = 0):
        self.lib = load_library()
        self.ctx = self.lib.hb_ctx_create(logn, logq, slots, logDelta, device)
        self.info = HbInfo()
        self.lib.hb_ctx_info(self.ctx, C.byref(self.info))
        self.n = self.info.n
        self.slots = slots
        self.L = self.info.L

    def close(self):
        if self.ctx:
            self.lib.hb_ctx_destroy(self.ctx)
            self.ctx = None

    # ---- context queries --------------------------------------------------------------------------
    def words(self, l: int) -> int:
        return self.lib.hb_ctx_words(self.ctx, l)

    def logq(self, l: int) -> int:
        return self.lib.hb_ctx_logq(self.ctx, l)

    def primes(self) -> np.ndarray:
        out = np.zeros(self.info.dimub, dtype=np.uint64)
        self.lib.hb_ctx_primes(self.ctx, _ptr(out))
        return out

    def twiddles(self, d: int, inverse: bool = False) -> np.ndarray:
        out = np.zeros(self.n, dtype=np.uint64)
        self.lib.hb_ctx_twiddles(self.ctx, d, int(inverse), _ptr(out))
        return out

    def sync(self):
        self.lib.hb_sync(self.ctx)

    def launches(self) -> int:
        return int(self.lib.hb_launch_count(self.ctx))

    # ---- device buffers -------------------------------------------------------------------------------
    def to_device(self, a: np.ndarray):
        a = np.ascontiguousarray(a)
        d = self.lib.hb_dev_alloc(self.ctx, a.nbytes)
        self.lib.hb_h2d(self.ctx, d, _ptr(a), a.nbytes)
        self.lib.hb_sync(self.ctx)
        return d

    def from_device(self, d, shape, dtype=np.uint64) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        self.lib.hb_d2h(self.ctx, _ptr(out), d, out.nbytes)
        self.lib.hb_sync(self.ctx)
        return out

    def dev_alloc(self, nbytes: int):
        return self.lib.hb_dev_alloc(self.ctx, nbytes)

    def dev_free(self, d):
        self.lib.hb_dev_free(self.ctx, d)

    # ---- containers -----------------------------------------------------------------------------------
    def ct(self, words: np.ndarray | None = None, l: int | None = None, nu=1.0, B=1.0, batch: int | None = None) -> Ciphertext:
        if words is not None:
            words = np.asarray(words, dtype=np.uint64)
            if words.ndim == 3:
                words = words[None]
            c = Ciphertext(self, words.shape[0])
            return c.upload(words, self.L if l is None else l, nu, B)
        return Ciphertext(self, batch or 1)

    def keys(self, arr: np.ndarray) -> EvalKeys:
        return EvalKeys(self, arr)

    def pt(self, words: np.ndarray, nu: float) -> HbPt:
        words = np.ascontiguousarray(words, dtype=np.uint64)
        assert words.shape == (self.info.words_top, self.n)
        p = HbPt()
        self.lib.hb_pt_alloc(self.ctx, C.byref(p))
        self.lib.hb_pt_upload(self.ctx, C.byref(p), _ptr(words), nu)
        return p

    def pt_free(self, p: HbPt):
        self.lib.hb_pt_free(self.ctx, C.byref(p))

    # ---- encode / decode (host) -----------------------------------------------------------------------
    def encode(self, m: np.ndarray, Delta: float | None = None) -> np.ndarray:
        m = np.ascontiguousarray(m, dtype=np.complex128)
        assert m.shape == (self.slots,)
        words = np.zeros((self.info.words_top, self.n), dtype=np.uint64)
        self.lib.hb_encode(self.ctx, _ptr(m), float(Delta or 2.0 ** self.info.logDelta), _ptr(words))
        return words

    def decode(self, words: np.ndarray, Delta: float) -> np.ndarray:
        words = np.ascontiguousarray(words, dtype=np.uint64)
        m = np.zeros(self.slots, dtype=np.complex128)
        self.lib.hb_decode(self.ctx, _ptr(words), words.shape[0], float(Delta), _ptr(m))
        return m

    def he_ecd(self, m: np.ndarray) -> HbPt:
        m = np.ascontiguousarray(m, dtype=np.complex128)
        p = HbPt()
        self.lib.hb_pt_alloc(self.ctx, C.byref(p))
        self.lib.hb_he_ecd(self.ctx, C.byref(p), _ptr(m))
        return p

    # ---- evaluation (names as in gpqhe.h) ---------------------------------------------------------------
    def _new_like(self, a: Ciphertext) -> Ciphertext:
        return Ciphertext(self, a.batch)

    def he_copy_ct(self, a: Ciphertext) -> Ciphertext:
        r = self._new_like(a)
        self.lib.hb_ct_copy(self.ctx, C.byref(r.h), C.byref(a.h))
        return r

    def he_add(self, a, b, out=None):
        r = out or self._new_like(a)
        self.lib.hb_he_add(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(b.h))
        return r

    def he_sub(self, a, b, out=None):
        r = out or self._new_like(a)
        self.lib.hb_he_sub(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(b.h))
        return r

    def he_neg(self, a):
        self.lib.hb_he_neg(self.ctx, C.byref(a.h))
        return a

    def he_rs(self, a):
        self.lib.hb_he_rs(self.ctx, C.byref(a.h))
        return a

    def he_moddown(self, a):
        self.lib.hb_he_moddown(self.ctx, C.byref(a.h))
        return a

    def he_mulpt(self, a, pt: HbPt, out=None):
        r = out or self._new_like(a)
        self.lib.hb_he_mulpt(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(pt))
        return r

    def he_mulpt_static(self, a, pt: HbPt, out):
        """hb_he_mulpt_static: he_mulpt without any allocation (scratch in the library's static device array);
        the caller supplies the output ciphertext."""
        self.lib.hb_he_mulpt_static(self.ctx, C.byref(out.h), C.byref(a.h), C.byref(pt))
        return out

    def he_mul(self, a, b, rlk: EvalKeys, out=None):
        r = out or self._new_like(a)
        self.lib.hb_he_mul(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(b.h), rlk.arr)
        return r

    def he_const_pt(self, num: complex, pt: HbPt | None = None) -> HbPt:
        """he_const_pt: writes coefficients 0 and N/2 only; a fresh handle is zero-filled."""
        if pt is None:
            pt = HbPt()
            self.lib.hb_pt_alloc(self.ctx, C.byref(pt))
        num = complex(num)
        self.lib.hb_he_const_pt(self.ctx, C.byref(pt), num.real, num.imag)
        return pt

    def he_addpt(self, a, pt: HbPt, out=None):
        r = out or self._new_like(a)
        self.lib.hb_he_addpt(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(pt))
        return r

    def he_subpt(self, a, pt: HbPt, out=None):
        r = out or self._new_like(a)
        self.lib.hb_he_subpt(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(pt))
        return r

    # ---- he-algo.c compositions (hb_algo.cpp) -------------------------------------------------------------
    def he_sum(self, a, rk: EvalKeys):
        r = self._new_like(a)
        self.lib.hb_he_sum(self.ctx, C.byref(r.h), C.byref(a.h), rk.arr)
        return r

    def he_idx(self, a, idx: int, rk: EvalKeys):
        r = self._new_like(a)
        self.lib.hb_he_idx(self.ctx, C.byref(r.h), C.byref(a.h), idx, rk.arr)
        return r

    def he_nrm2(self, a, rlk: EvalKeys, ck: EvalKeys, rk: EvalKeys):
        r = self._new_like(a)
        self.lib.hb_he_nrm2(self.ctx, C.byref(r.h), C.byref(a.h), rlk.arr, ck.arr, rk.arr)
        return r

    def he_inv(self, a, rlk: EvalKeys, iters: int):
        r = self._new_like(a)
        self.lib.hb_he_inv(self.ctx, C.byref(r.h), C.byref(a.h), rlk.arr, iters)
        return r

    def he_sqrt(self, a, rlk: EvalKeys, iters: int):
        r = self._new_like(a)
        self.lib.hb_he_sqrt(self.ctx, C.byref(r.h), C.byref(a.h), rlk.arr, iters)
        return r

    def he_sigmoid(self, a, rlk: EvalKeys):
        r = self._new_like(a)
        self.lib.hb_he_sigmoid(self.ctx, C.byref(r.h), C.byref(a.h), rlk.arr)
        return r

    def he_log(self, a, rlk: EvalKeys):
        r = self._new_like(a)
        self.lib.hb_he_log(self.ctx, C.byref(r.h), C.byref(a.h), rlk.arr)
        return r

    def he_exp(self, coef: complex, a, rlk: EvalKeys, iters: int):
        r = self._new_like(a)
        coef = complex(coef)
        self.lib.hb_he_exp(self.ctx, C.byref(r.h), coef.real, coef.imag, C.byref(a.h), rlk.arr, iters)
        return r

    def he_cmp(self, a, b, rlk: EvalKeys, iters: int, alpha: int):
        r = self._new_like(a)
        self.lib.hb_he_cmp(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(b.h), rlk.arr, iters, alpha)
        return r

    def he_cmppt(self, a, pt: HbPt, rlk: EvalKeys, iters: int, alpha: int):
        r = self._new_like(a)
        self.lib.hb_he_cmppt(self.ctx, C.byref(r.h), C.byref(a.h), C.byref(pt), rlk.arr, iters, alpha)
        return r

    def he_rot(self, a, rot: int, rk: EvalKeys):
        self.lib.hb_he_rot(self.ctx, C.byref(a.h), rot, rk.arr)
        return a

    def he_conj(self, a, ck: EvalKeys):
        self.lib.hb_he_conj(self.ctx, C.byref(a.h), ck.arr)
        return a

    def he_gemv(self, A: np.ndarray, a, rk: EvalKeys, out=None):
        A = np.ascontiguousarray(A, dtype=np.complex128)
        assert A.shape == (self.slots, self.slots)
        r = out or self._new_like(a)
        self.lib.hb_he_gemv(self.ctx, C.byref(r.h), _ptr(A), C.byref(a.h), rk.arr)
        return r

    def ctr_hempc_eval(self, Ka, Kb, xhat, xr, uhat, ur, rk: EvalKeys, out=None):
        Ka = np.ascontiguousarray(Ka, dtype=np.complex128)
        Kb = np.ascontiguousarray(Kb, dtype=np.complex128)
        r = out or self._new_like(xhat)
        self.lib.hb_ctr_hempc_eval(self.ctx, C.byref(r.h), _ptr(Ka), _ptr(Kb), C.byref(xhat.h), C.byref(xr.h),
                                   C.byref(uhat.h), C.byref(ur.h), rk.arr)
        return r

    def ctr_hempc_eval_host(self, Ka, Kb, xhat, xr, uhat, ur, rk: EvalKeys, l: int, out: np.ndarray | None = None):
        """Host arrays in ([batch][2][W_l][N]) -> host array out ([batch][2][W_{l-1}][N]); H2D and D2H inside."""
        Ka = np.ascontiguousarray(Ka, dtype=np.complex128)
        Kb = np.ascontiguousarray(Kb, dtype=np.complex128)
        batch = xhat.shape[0]
        if out is None:
            out = np.empty((batch, 2, self.words(l - 1), self.n), dtype=np.uint64)
        self.lib.hb_ctr_hempc_eval_host(self.ctx, batch, l, _ptr(Ka), _ptr(Kb), _ptr(xhat), _ptr(xr), _ptr(uhat),
                                        _ptr(ur), rk.arr, _ptr(out))
        return out

    def ctr_hempc_submit_host(self, Ka, Kb, xhat, xr, uhat, ur, rk: EvalKeys, l: int, out: np.ndarray) -> int:
        """Streaming form (hb_ctr_hempc_submit_host): returns a ticket; ``out`` is complete after ``ctr_hempc_wait(ticket)``.
        The arrays must stay alive (and should be pinned) until then."""
        Ka = np.ascontiguousarray(Ka, dtype=np.complex128)
        Kb = np.ascontiguousarray(Kb, dtype=np.complex128)
        for a in (xhat, xr, uhat, ur, out):
            assert a.flags["C_CONTIGUOUS"] and a.dtype == np.uint64
        return int(self.lib.hb_ctr_hempc_submit_host(self.ctx, xhat.shape[0], l, _ptr(Ka), _ptr(Kb), _ptr(xhat), _ptr(xr), _ptr(uhat),
                                                     _ptr(ur), rk.arr, _ptr(out)))

    def ctr_hempc_wait(self, ticket: int):
        self.lib.hb_ctr_hempc_wait(self.ctx, ticket)

    # ---- client side on the device (he-encrypt.c) ---------------------------------------------------------
    def pk(self, p0_words: np.ndarray, p1_words: np.ndarray) -> HbEvk:
        """he_pk (gpqhe.h:72-76) as words[W_L][N] each -> device transform limbs."""
        p0 = np.ascontiguousarray(p0_words, dtype=np.uint64)
        p1 = np.ascontiguousarray(p1_words, dtype=np.uint64)
        assert p0.shape == p1.shape == (self.info.words_top, self.n)
        k = HbEvk()
        self.lib.hb_pk_upload(self.ctx, C.byref(k), _ptr(p0), _ptr(p1))
        return k

    def sk(self, s: np.ndarray) -> HbSk:
        s = np.ascontiguousarray(s, dtype=np.int64)
        assert s.shape == (self.n,)
        k = HbSk()
        self.lib.hb_sk_upload(self.ctx, C.byref(k), _ptr(s))
        return k

    # ---- wire format (hb_wire.cpp) -----------------------------------------------------------------------------
    def ct_serialize(self, a: Ciphertext) -> bytes:
        n = int(self.lib.hb_wire_ct_bytes(self.ctx, a.l, a.batch))
        buf = C.create_string_buffer(n)
        got = int(self.lib.hb_ct_serialize(self.ctx, C.byref(a.h), buf, n))
        assert got == n
        return buf.raw

    def ct_deserialize(self, msg: bytes, batch: int) -> Ciphertext:
        r = Ciphertext(self, batch)
        self.lib.hb_ct_deserialize(self.ctx, C.byref(r.h), msg, len(msg))
        return r

    def evk_serialize(self, key: np.ndarray) -> bytes:
        key = np.ascontiguousarray(key, dtype=np.uint64)
        assert key.shape == (2, self.info.dimevk, self.n)
        n = int(self.lib.hb_wire_evk_bytes(self.ctx))
        buf = C.create_string_buffer(n)
        assert int(self.lib.hb_evk_serialize(self.ctx, _ptr(key[0]), _ptr(key[1]), buf, n)) == n
        return buf.raw

    def evk_deserialize(self, msg: bytes) -> HbEvk:
        k = HbEvk()
        self.lib.hb_evk_deserialize(self.ctx, C.byref(k), msg, len(msg))
        return k

    def genswk_words(self) -> int:
        return int(self.lib.hb_genswk_words(self.ctx))

    def he_genswk(self, a_words: np.ndarray, e: np.ndarray, sp: np.ndarray, sk: HbSk) -> np.ndarray:
        """he_genswk with the caller's samples, nkeys keys at once: a_words[nkeys][genswk_words][N] (sample_uniform),
        e, sp int64[nkeys][N] -> uint64[nkeys][2][dimevk][N], the layout ``Engine.keys`` takes."""
        a_words = np.ascontiguousarray(a_words, dtype=np.uint64)
        nkeys = a_words.shape[0]
        assert a_words.shape == (nkeys, self.genswk_words(), self.n)
        e = np.ascontiguousarray(e, dtype=np.int64)
        sp = np.ascontiguousarray(sp, dtype=np.int64)
        assert e.shape == sp.shape == (nkeys, self.n)
        p0 = np.empty((nkeys, self.info.dimevk, self.n), dtype=np.uint64)
        p1 = np.empty_like(p0)
        self.lib.hb_he_genswk(self.ctx, nkeys, _ptr(a_words), _ptr(e), _ptr(sp), C.byref(sk), _ptr(p0), _ptr(p1))
        return np.stack([p0, p1], axis=1)

    def he_enc_pk(self, pt_words: np.ndarray, pk: HbEvk, v: np.ndarray, e0: np.ndarray, e1: np.ndarray, pt_nu=None) -> Ciphertext:
        """he_enc_pk with the caller's samples: pt_words[batch][W_L][N], v/e0/e1 int64[batch][N]."""
        pt_words = np.ascontiguousarray(pt_words, dtype=np.uint64)
        batch = pt_words.shape[0]
        assert pt_words.shape == (batch, self.info.words_top, self.n)
        arrs = [np.ascontiguousarray(a, dtype=np.int64) for a in (v, e0, e1)]
        assert all(a.shape == (batch, self.n) for a in arrs)
        ct = Ciphertext(self, batch)
        self.lib.hb_he_enc_pk(self.ctx, C.byref(ct.h), _ptr(pt_words), float(pt_nu or 2.0 ** self.info.logDelta), C.byref(pk),
                              _ptr(arrs[0]), _ptr(arrs[1]), _ptr(arrs[2]))
        return ct

    def he_dec(self, ct: Ciphertext, sk: HbSk) -> np.ndarray:
        """he_dec -> plaintext words[batch][W_l][N] on the host."""
        out = np.empty((ct.batch, self.words(ct.l), self.n), dtype=np.uint64)
        self.lib.hb_he_dec(self.ctx, _ptr(out), C.byref(ct.h), C.byref(sk))
        return out

    # ---- limb level -----------------------------------------------------------------------------------------
    def ntt(self, limbs: np.ndarray, d0: int = 0, D: int = 1, inverse=False, strict=False) -> np.ndarray:
        limbs = np.ascontiguousarray(limbs, dtype=np.uint64)
        flat = limbs.reshape(-1, self.n)
        d = self.to_device(flat)
        if inverse:
            self.lib.hb_ntt_inverse(self.ctx, d, d, flat.shape[0], d0, D)
        else:
            self.lib.hb_ntt_forward(self.ctx, d, d, flat.shape[0], d0, D, int(strict))
        out = self.from_device(d, flat.shape).reshape(limbs.shape)
        self.dev_free(d)
        return out

    def poly_ntt(self, polys: np.ndarray, D: int) -> np.ndarray:
        polys = np.ascontiguousarray(polys, dtype=np.uint64)        # [npoly][W][N]
        npoly, W, _ = polys.shape
        d = self.to_device(polys)
        o = self.dev_alloc(npoly * D * self.n * 8)
        self.lib.hb_poly_ntt(self.ctx, d, W, npoly, D, o)
        out = self.from_device(o, (npoly, D, self.n))
        self.dev_free(d)
        self.dev_free(o)
        return out

    def rns2words(self, limbs: np.ndarray, k: int, Wout: int) -> np.ndarray:
        limbs = np.ascontiguousarray(limbs, dtype=np.uint64)        # [npoly][D][N]
        npoly, D, _ = limbs.shape
        d = self.to_device(limbs)
        o = self.dev_alloc(npoly * Wout * self.n * 8)
        self.lib.hb_rns2words(self.ctx, d, npoly, D, k, Wout, o)
        out = self.from_device(o, (npoly, Wout, self.n))
        self.dev_free(d)
        self.dev_free(o)
        return out

    def poly_rot(self, polys: np.ndarray, rot: int) -> np.ndarray:
        polys = np.ascontiguousarray(polys, dtype=np.uint64)
        npoly, W, _ = polys.shape
        d = self.to_device(polys)
        o = self.dev_alloc(polys.nbytes)
        self.lib.hb_poly_rot(self.ctx, d, W, npoly, rot, o)
        out = self.from_device(o, polys.shape)
        self.dev_free(d)
        self.dev_free(o)
        return out
