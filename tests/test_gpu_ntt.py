"""GPU parity of the stand-alone transforms (config B): hb_ntt_forward / hb_ntt_inverse against ntt() / invntt()
of the compiled reference (GPQHE/src/ntt.c:37-73), bit for bit, for every ring size class (1, 2, 4, 8 CTAs per limb).
Kept in its own module because the reference has ONE process-global context (precomp.c:41-47)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("logn", [7, 10, 11, 12, 13, 14, 15, 16])
def test_ntt_bit_exact(ref_available, logn):
    from oracle.ref import Ref
    from hectr_b200 import Engine
    logq = {7: 61, 10: 27, 11: 54}.get(logn, 109)
    slots, logD = (16, 50) if logn > 11 else (16, 20)
    ref = Ref.get(logn, logq, slots, logD)
    eng = Engine(logn, logq, slots, logD)
    try:
        n = 1 << logn
        rng = np.random.default_rng(1234 + logn)
        primes = ref.primes()
        D = min(3, len(primes))
        limbs = np.stack([rng.integers(0, primes[d], size=n, dtype=np.uint64) for d in range(D)] * 2)   # 2*D limbs
        # adversarial limbs: zeros, all p-1, and p itself as an input
        adv = np.zeros((3, n), dtype=np.uint64)
        adv[1, :] = primes[0] - 1
        adv[2, ::2] = primes[0]
        want = np.stack([ref.ntt(limbs[j], j % D) for j in range(limbs.shape[0])])
        got_strict = eng.ntt(limbs, 0, D, strict=True)
        assert np.array_equal(got_strict, want), "strict forward NTT differs from ntt.c"
        got = eng.ntt(limbs, 0, D, strict=False)
        canon = np.stack([want[j] % np.uint64(primes[j % D]) for j in range(limbs.shape[0])])
        assert np.array_equal(got, canon), "lazy forward NTT differs from canonicalised ntt.c"
        want_adv = np.stack([ref.ntt(adv[j], 0) for j in range(3)])
        assert np.array_equal(eng.ntt(adv, 0, 1, strict=True), want_adv)
        assert np.array_equal(eng.ntt(adv, 0, 1, strict=False), want_adv % np.uint64(primes[0]))
        # inverse: inputs as the reference would feed them (forward outputs, in [0,p])
        want_inv = np.stack([ref.ntt(want[j], j % D, inverse=True) for j in range(limbs.shape[0])])
        assert np.array_equal(eng.ntt(want, 0, D, inverse=True), want_inv)
        assert np.array_equal(want_inv, limbs)     # round trip (size-independent property)
        assert np.array_equal(eng.ntt(want_adv, 0, 1, inverse=True), np.stack([ref.ntt(want_adv[j], 0, inverse=True) for j in range(3)]))
    finally:
        eng.close()


