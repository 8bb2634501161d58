"""Writes tests/golden/config_c_chain.json: SHA-256 of every ciphertext the UNMODIFIED reference (oracle/_ref) produces along
BASELINE configs[2] -- N = 2^15, q = 2^881, Delta = 2^50: from reference-encrypted m1, m2
(GPQHE/tests/gpqhe.c:200-203) at level 17 down to level 1:  he_rot(a, 1), he_mul(a, b) + relinearise, he_rs, he_moddown(b).
Deterministic (SUPERCOP stream, GPQHE/src/rng.c:32-78).  Run where /root/reference exists:  python tests/golden/make_config_c.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

PARAMS = (15, 881, 2, 50)          # slots = 2: he_genrk makes `slots` keys of 23.5 MiB each; he_rot(., 1) needs rk[1] only


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()


def setup(ref):
    S = ref.slots
    pk, sk = ref.keypair()
    rlk = ref.genrlk(sk)
    rk = ref.genrk(sk)
    m1 = np.array([(1 - i / S) * (1 + 1j) for i in range(S)], dtype=np.complex128)
    m2 = np.array([(0.99 - i / S) * (1 + 1j) for i in range(S)], dtype=np.complex128)
    a = ref.enc_pk(ref.encode(m1), pk)
    b = ref.enc_pk(ref.encode(m2), pk)
    return pk, sk, rlk, rk, a, b, m1, m2


def main():
    from oracle.ref import Ref
    ref = Ref.get(*PARAMS)
    pk, sk, rlk, rk, a, b, m1, m2 = setup(ref)
    out = {"params": PARAMS, "inputs": {"a": sha(ref.ct_to_words(a)), "b": sha(ref.ct_to_words(b)),
                                        "rlk": sha(ref.evk_to_array(rlk)), "rk1": sha(ref.evk_to_array(rk[1]))},
           "levels": {}}
    for l in range(ref.L, 0, -1):
        rec = {}
        r = ref.rot(ref.copy_ct(a), 1, rk)
        rec["rot"] = sha(ref.ct_to_words(r))
        ref.free_ct(r)
        p = ref.mul(a, b, rlk)
        rec["mul"] = sha(ref.ct_to_words(p))
        ref.rs(p)
        rec["rs"] = sha(ref.ct_to_words(p))
        ref.moddown(b)
        rec["moddown"] = sha(ref.ct_to_words(b))
        ref.free_ct(a)
        a = p
        out["levels"][str(l)] = rec
        print(l, rec["mul"][:12], flush=True)
    json.dump(out, open(os.path.join(ROOT, "tests", "golden", "config_c_chain.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
