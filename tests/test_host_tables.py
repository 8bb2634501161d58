"""CPU tests of the engine's host logic (no GPU): the table builder, limb-count formulas and the host
encoder of libhectr_b200.so against the compiled reference (oracle/_ref)."""
import ctypes as C

import numpy as np
import pytest

from hectr_b200 import _lib


def _tables(logn, logq, slots, logD):
    lib = _lib.load_library()
    t = lib.hb_tables_create(logn, logq, slots, logD)
    assert t
    info = _lib.HbInfo()
    lib.hb_tables_info(t, C.byref(info))
    return lib, t, info


def test_library_exports_every_declared_symbol():
    lib = _lib.load_library()
    declared = _lib.header_symbols()
    assert len(declared) >= 50
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    assert sorted(_lib.PROTOTYPES) == declared, "ctypes prototypes out of sync with include/hectr_b200.h"


@pytest.mark.parametrize("params", [(12, 109, 16, 50), (7, 61, 16, 50), (13, 218, 32, 50), (16, 109, 16, 50)])
def test_tables_match_reference(ref_available, params):
    from oracle.ref import Ref
    ref = Ref.get(*params)
    lib, t, info = _tables(*params)
    try:
        assert (info.L, info.dim, info.dimevk, info.dimub) == (ref.L, ref.dim, ref.dimevk, ref.dimub)
        assert info.nbitsP == ref.mpi_int(ref.hectx.P).bit_length()
        assert info.nbitsPqL == ref.mpi_int(ref.hectx.PqL).bit_length()
        primes = np.zeros(info.dimub, dtype=np.uint64)
        lib.hb_tables_primes(t, primes.ctypes.data_as(C.c_void_p))
        assert list(map(int, primes)) == ref.primes()
        n = info.n
        tw = np.zeros(n, dtype=np.uint64)
        for d in sorted({0, 1, info.dimub - 1}):
            p = int(primes[d])
            for inv in (0, 1):
                lib.hb_tables_twiddles(t, d, inv, tw.ctypes.data_as(C.c_void_p))
                mont = ref.zetas(d, inv=bool(inv))          # zeta * 2^64 mod p  (precomp.c:258-259)
                idx = np.random.default_rng(d).integers(0, n, size=64)
                for i in list(idx) + [0, 1, n - 1]:
                    assert (int(tw[i]) << 64) % p == int(mont[i]), (d, inv, i)
        # CRT constants (precomp.c:266-293) for a few limb counts
        WC = info.words_top
        for D in sorted({1, 2, info.dim, info.dimevk}):
            inv = np.zeros(D, dtype=np.uint64)
            phat = np.zeros((D, WC), dtype=np.uint64)
            Plow = np.zeros(WC, dtype=np.uint64)
            lib.hb_tables_crt(t, D, inv.ctypes.data_as(C.c_void_p), phat.ctypes.data_as(C.c_void_p), Plow.ctypes.data_as(C.c_void_p))
            node = ref.node(D - 1)
            P = ref.mpi_int(node.P)
            mask = (1 << (64 * WC)) - 1
            assert sum(int(Plow[w]) << (64 * w) for w in range(WC)) == P & mask
            for d in range(D):
                assert int(inv[d]) == int(node.phat_invmp[d])
                assert sum(int(phat[d, w]) << (64 * w) for w in range(WC)) == ref.mpi_int(node.phat[d]) & mask
    finally:
        lib.hb_tables_destroy(t)


def test_limb_count_formulas(ref_available):
    """he-automorphism.c:52, he-mult.c:99,168 -- compared with what the formulas give on the reference's own context."""
    from oracle.ref import Ref
    params = (12, 109, 16, 50)
    ref = Ref.get(*params)
    lib, t, info = _tables(*params)
    nbPqL = ref.mpi_int(ref.hectx.PqL).bit_length()
    for l in range(ref.L + 1):
        nbq = ref.q_level(l).bit_length()
        assert lib.hb_tables_dim(t, 0, l, 0.0) == (nbq + nbPqL + 12) // 59 + 1
        assert lib.hb_tables_dim(t, 1, l, 0.0) == (2 * nbq + 12) // 59 + 1
        for nu in (2.0 ** 50, 3.7e15, 1.0):
            assert lib.hb_tables_dim(t, 2, l, nu) == int((nbq + np.log2(nu) + 12) / 59 + 1)
    assert (lib.hb_tables_dim(t, 0, 2, 0), lib.hb_tables_dim(t, 2, 2, 2.0 ** 50), lib.hb_tables_dim(t, 1, 2, 0)) == (7, 3, 4)
    lib.hb_tables_destroy(t)


@pytest.mark.parametrize("params", [(12, 109, 16, 50), (13, 218, 64, 50), (12, 109, 128, 50)])
def test_host_encoder_bit_exact(ref_available, params):
    """hb_tables_encode / decode against he_ecd / he_dcd (he-encode.c:53-75,110-125)."""
    from oracle.ref import Ref
    ref = Ref.get(*params)
    lib, t, info = _tables(*params)
    rng = np.random.default_rng(5)
    S, WC, n = params[2], info.words_top, info.n
    for trial in range(4):
        m = (rng.standard_normal(S) + 1j * rng.standard_normal(S)) * (10.0 ** (trial - 1))
        if trial == 3:
            m[:] = 0
            m[:3] = [1.0, -0.5, 0.25j]
        pt = ref.encode(m)
        want = ref.poly_to_words(pt.m, WC)
        got = np.zeros((WC, n), dtype=np.uint64)
        mm = np.ascontiguousarray(m)
        lib.hb_tables_encode(t, mm.ctypes.data_as(C.c_void_p), 2.0 ** params[3], got.ctypes.data_as(C.c_void_p))
        assert np.array_equal(got, want)
        back_ref = ref.decode(pt)
        back = np.zeros(S, dtype=np.complex128)
        lib.hb_tables_decode(t, got.ctypes.data_as(C.c_void_p), WC, 2.0 ** params[3], back.ctypes.data_as(C.c_void_p))
        assert np.array_equal(back, back_ref)
        assert np.abs(back - m).max() < 1e-9 * max(1.0, np.abs(m).max())
    lib.hb_tables_destroy(t)


def test_rejects_what_the_reference_rejects():
    lib = _lib.load_library()
    assert not lib.hb_tables_create(12, 110, 16, 50)      # log q above the security bound (precomp.c:344-350)
    assert not lib.hb_tables_create(12, 109, 12, 50)      # slots not a power of two (precomp.c:434-439)
    assert not lib.hb_tables_create(12, 109, 4096, 50)    # slots > n/2 (precomp.c:440-445)
    assert not lib.hb_tables_create(12, 109, 16, 10)      # Delta below the fresh-noise bound (precomp.c:452)
