"""The additive wire format (hb_wire.cpp): round trips through bytes, header fields at their documented
offsets, and loud failure on a damaged message.  The reference has no serialisation to compare with."""
import ctypes as C
import struct
import subprocess
import sys
import os

import numpy as np
import pytest

from conftest import random_ct_words

pytestmark = pytest.mark.gpu
PARAMS = (12, 109, 16, 50)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def eng():
    from hectr_b200 import Engine
    e = Engine(*PARAMS)
    yield e
    e.close()


def test_ct_round_trip_and_header(eng):
    rng = np.random.default_rng(5)
    for l, batch in ((2, 3), (1, 1)):
        W = eng.words(l)
        w = random_ct_words(rng, batch, W, eng.n, eng.logq(l))
        a = eng.ct(w, l, nu=2.0 ** 50, B=123.5)
        msg = eng.ct_serialize(a)
        assert len(msg) == 72 + w.nbytes
        assert msg[:4] == b"HBCT"
        version, logn, logq, logDelta, slots, level, b, words, reserved = struct.unpack_from("<9I", msg, 4)
        assert (version, logn, logq, logDelta, slots, level, b, words, reserved) == (1, 12, 109, 50, 16, l, batch, W, 0)
        nu, B, payload = struct.unpack_from("<ddQ", msg, 40)
        assert (nu, B, payload) == (2.0 ** 50, 123.5, w.nbytes)
        assert msg[72:] == w.tobytes()
        back = eng.ct_deserialize(msg, batch)
        assert (back.l, back.nu, back.B) == (l, 2.0 ** 50, 123.5)
        assert np.array_equal(back.download(), w)


def test_evk_round_trip_gives_the_same_rotation(eng):
    rng = np.random.default_rng(6)
    primes = eng.primes()
    key = np.empty((2, eng.info.dimevk, eng.n), dtype=np.uint64)
    for d in range(eng.info.dimevk):
        key[:, d, :] = rng.integers(0, int(primes[d]), size=(2, eng.n), dtype=np.uint64)
    msg = eng.evk_serialize(key)
    assert msg[:4] == b"HBEK" and len(msg) == 72 + key.nbytes and msg[72:] == key.tobytes()
    from hectr_b200._lib import HbEvk
    arr = (HbEvk * 1)()
    arr[0] = eng.evk_deserialize(msg)
    direct = eng.keys(key)
    w = random_ct_words(rng, 2, eng.words(2), eng.n, eng.logq(2))

    class _K:            # the rotation entry point takes the array of keys
        pass
    k = _K()
    k.arr = arr
    got = eng.he_rot(eng.ct(w, 2), 0, k).download()
    want = eng.he_rot(eng.ct(w, 2), 0, direct).download()
    assert np.array_equal(got, want)
    eng.lib.hb_evk_free(eng.ctx, C.byref(arr[0]))
    direct.free()


CHILD = r"""
import sys
sys.path.insert(0, %(root)r)
sys.path.insert(0, %(root)r + "/tests")
import numpy as np
from conftest import random_ct_words
from hectr_b200 import Engine
eng = Engine(12, 109, 16, 50)
w = random_ct_words(np.random.default_rng(1), 1, eng.words(2), eng.n, eng.logq(2))
msg = bytearray(eng.ct_serialize(eng.ct(w, 2)))
mode = sys.argv[1]
if mode == "flip":
    msg[5000] ^= 1
elif mode == "params":
    msg[12] ^= 1            # logq
elif mode == "short":
    msg = msg[:-8]
eng.ct_deserialize(bytes(msg), 1)
print("accepted")
"""


@pytest.mark.parametrize("mode,text", [("flip", "checksum mismatch"), ("params", "parameters differ"), ("short", "truncated")])
def test_damaged_message_aborts(mode, text):
    r = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}, mode], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "accepted" not in r.stdout
    assert text in r.stderr, r.stderr[-500:]
