"""CPU-side checks of the reference-side binding (integration/gpqhe_b200_shim.c) and of the bench's key helpers.  No GPU work:
the shim library is only inspected, never called."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "integration", "_build")

# the evaluation half of GPQHE/src/gpqhe.h:99-175 that HECTR reaches (src/hempc.c:253-266, he-algo.c:47-93), plus the context pair
SHIM_SYMBOLS = ["hectx_init", "hectx_exit", "he_add", "he_sub", "he_neg", "he_addpt", "he_subpt", "he_rs", "he_moddown", "he_mulpt",
                "he_mul", "he_rot", "he_conj", "he_gemv", "he_enc_pk", "he_dec"]


def _defined(path):
    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if " T " in line}


def test_shim_exports_the_reference_names():
    lib = os.path.join(BUILD, "libgpqhe_b200.so")
    if not os.path.exists(lib):
        pytest.skip("integration/_build not built (integration/build.sh needs /root/reference)")
    have = _defined(lib)
    assert set(SHIM_SYMBOLS) <= have, sorted(set(SHIM_SYMBOLS) - have)
    # everything else must come from the reference: the shim defines no key generation, encoding, sampling or allocation
    for name in ("he_keypair", "he_genrk", "he_ecd", "he_dcd", "he_alloc_ct", "he_copy_ct", "sample_zo", "poly_mul"):
        assert name not in have, name


def test_driver_links_the_shim_before_the_reference():
    exe = os.path.join(BUILD, "test-hectr-b200")
    if not os.path.exists(exe):
        pytest.skip("integration/_build not built")
    needed = [l.split("[")[1].rstrip("]") for l in subprocess.run(["readelf", "-d", exe], capture_output=True, text=True, check=True).stdout.splitlines()
              if "(NEEDED)" in l]
    assert "libgpqhe_b200.so" in needed and "libhectr.so" in needed
    assert needed.index("libgpqhe_b200.so") < needed.index("libhectr.so")       # symbol lookup follows this order
    plain = [l for l in subprocess.run(["readelf", "-d", os.path.join(BUILD, "test-hectr")], capture_output=True, text=True, check=True).stdout.splitlines()
             if "(NEEDED)" in l]
    assert not any("gpqhe_b200" in l for l in plain)                             # the reference run stays pure


def test_bench_rotation_of_the_secret_matches_poly_rot(ref_available):
    """bench.py makes its rotation keys with hb_he_genswk from sp = poly_rot(sk, r): its numpy poly_rot must be the reference's
    (GPQHE/src/poly.c:263-283)."""
    import bench
    from oracle.ref import Ref
    ref = Ref.get(7, 150, 8, 40)
    rng = np.random.default_rng(5)
    a = rng.integers(-3, 4, size=ref.n).astype(np.int64)
    for r in (0, 1, 5, 7):
        want = ref.poly_rot(a.view(np.uint64)[None], r)[0].view(np.int64)
        assert np.array_equal(bench._poly_rot_i64(a, r), want), r
