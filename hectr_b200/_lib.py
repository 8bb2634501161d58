"""ctypes binding of ``include/hectr_b200.h`` (one prototype per exported symbol)."""
from __future__ import annotations

import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HB_LIB") or os.path.join(HERE, "libhectr_b200.so")   # HB_LIB: development variants
HEADER_PATH = os.path.join(HERE, os.pardir, "include", "hectr_b200.h")

u64p = C.POINTER(C.c_uint64)
f64p = C.POINTER(C.c_double)


class HbInfo(C.Structure):
    _fields_ = [(n, C.c_uint) for n in ("logn", "n", "logq", "logDelta", "slots", "L", "dim", "dimevk", "dimub",
                                         "words_top", "nbitsP", "nbitsPqL")] + [("device", C.c_int), ("sm_count", C.c_int)]


class HbCt(C.Structure):
    _fields_ = [("l", C.c_uint), ("nu", C.c_double), ("B", C.c_double), ("batch", C.c_uint), ("words", C.c_void_p)]


class HbPt(C.Structure):
    _fields_ = [("nu", C.c_double), ("words", C.c_void_p)]


class HbEvk(C.Structure):
    _fields_ = [("p", C.c_void_p)]


class HbSk(C.Structure):
    _fields_ = [("hat", C.c_void_p), ("scaled", C.c_void_p * 64)]


ctx_p = C.c_void_p
ct_p = C.POINTER(HbCt)
pt_p = C.POINTER(HbPt)
evk_p = C.POINTER(HbEvk)
vp = C.c_void_p

PROTOTYPES = {
    "hb_ctx_create": (ctx_p, [C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int]),
    "hb_ctx_destroy": (None, [ctx_p]),
    "hb_ctx_info": (None, [ctx_p, C.POINTER(HbInfo)]),
    "hb_ctx_words": (C.c_uint, [ctx_p, C.c_uint]),
    "hb_ctx_logq": (C.c_uint, [ctx_p, C.c_uint]),
    "hb_ctx_primes": (None, [ctx_p, vp]),
    "hb_ctx_twiddles": (None, [ctx_p, C.c_uint, C.c_int, vp]),
    "hb_dim_keyswitch": (C.c_uint, [ctx_p, C.c_uint]),
    "hb_dim_mul": (C.c_uint, [ctx_p, C.c_uint]),
    "hb_dim_mulpt": (C.c_uint, [ctx_p, C.c_uint, C.c_double]),
    "hb_ctx_set_stream": (None, [ctx_p, vp]),
    "hb_sync": (None, [ctx_p]),
    "hb_timer_start": (None, [ctx_p]),
    "hb_timer_stop_ms": (C.c_float, [ctx_p]),
    "hb_launch_count": (C.c_uint64, [ctx_p]),
    "hb_prof_enable": (None, [ctx_p, C.c_int]),
    "hb_prof_reset": (None, [ctx_p]),
    "hb_prof_report": (C.c_size_t, [ctx_p, C.c_char_p, C.c_size_t]),
    "hb_ct_export": (None, [ctx_p, ct_p, vp]),
    "hb_ct_import": (None, [ctx_p, ct_p, vp, C.c_uint, C.c_double, C.c_double]),
    "hb_tables_create": (vp, [C.c_uint, C.c_uint, C.c_uint, C.c_uint]),
    "hb_tables_destroy": (None, [vp]),
    "hb_tables_info": (None, [vp, C.POINTER(HbInfo)]),
    "hb_tables_primes": (None, [vp, vp]),
    "hb_tables_twiddles": (None, [vp, C.c_uint, C.c_int, vp]),
    "hb_tables_crt": (None, [vp, C.c_uint, vp, vp, vp]),
    "hb_tables_dim": (C.c_uint, [vp, C.c_int, C.c_uint, C.c_double]),
    "hb_tables_encode": (None, [vp, vp, C.c_double, vp]),
    "hb_tables_decode": (None, [vp, vp, C.c_uint, C.c_double, vp]),
    "hb_dev_alloc": (vp, [ctx_p, C.c_size_t]),
    "hb_dev_free": (None, [ctx_p, vp]),
    "hb_host_alloc": (vp, [ctx_p, C.c_size_t]),
    "hb_host_free": (None, [ctx_p, vp]),
    "hb_h2d": (None, [ctx_p, vp, vp, C.c_size_t]),
    "hb_d2h": (None, [ctx_p, vp, vp, C.c_size_t]),
    "hb_dev_fill": (None, [ctx_p, vp, C.c_int, C.c_size_t]),
    "hb_ct_alloc": (None, [ctx_p, ct_p, C.c_uint]),
    "hb_ct_free": (None, [ctx_p, ct_p]),
    "hb_ct_copy": (None, [ctx_p, ct_p, ct_p]),
    "hb_ct_upload": (None, [ctx_p, ct_p, vp, C.c_uint, C.c_double, C.c_double]),
    "hb_ct_download": (None, [ctx_p, ct_p, vp]),
    "hb_pt_alloc": (None, [ctx_p, pt_p]),
    "hb_pt_free": (None, [ctx_p, pt_p]),
    "hb_pt_upload": (None, [ctx_p, pt_p, vp, C.c_double]),
    "hb_evk_upload": (None, [ctx_p, evk_p, vp, vp]),
    "hb_evk_free": (None, [ctx_p, evk_p]),
    "hb_encode": (None, [ctx_p, vp, C.c_double, vp]),
    "hb_decode": (None, [ctx_p, vp, C.c_uint, C.c_double, vp]),
    "hb_he_ecd": (None, [ctx_p, pt_p, vp]),
    "hb_he_add": (None, [ctx_p, ct_p, ct_p, ct_p]),
    "hb_he_sub": (None, [ctx_p, ct_p, ct_p, ct_p]),
    "hb_he_neg": (None, [ctx_p, ct_p]),
    "hb_he_addpt": (None, [ctx_p, ct_p, ct_p, pt_p]),
    "hb_he_subpt": (None, [ctx_p, ct_p, ct_p, pt_p]),
    "hb_he_const_pt": (None, [ctx_p, pt_p, C.c_double, C.c_double]),
    "hb_he_sum": (None, [ctx_p, ct_p, ct_p, evk_p]),
    "hb_he_idx": (None, [ctx_p, ct_p, ct_p, C.c_uint, evk_p]),
    "hb_he_nrm2": (None, [ctx_p, ct_p, ct_p, evk_p, evk_p, evk_p]),
    "hb_he_inv": (None, [ctx_p, ct_p, ct_p, evk_p, C.c_uint]),
    "hb_he_sqrt": (None, [ctx_p, ct_p, ct_p, evk_p, C.c_uint]),
    "hb_he_sigmoid": (None, [ctx_p, ct_p, ct_p, evk_p]),
    "hb_he_log": (None, [ctx_p, ct_p, ct_p, evk_p]),
    "hb_he_exp": (None, [ctx_p, ct_p, C.c_double, C.c_double, ct_p, evk_p, C.c_uint]),
    "hb_he_cmp": (None, [ctx_p, ct_p, ct_p, ct_p, evk_p, C.c_uint, C.c_uint]),
    "hb_he_cmppt": (None, [ctx_p, ct_p, ct_p, pt_p, evk_p, C.c_uint, C.c_uint]),
    "hb_he_rs": (None, [ctx_p, ct_p]),
    "hb_he_moddown": (None, [ctx_p, ct_p]),
    "hb_he_mulpt": (None, [ctx_p, ct_p, ct_p, pt_p]),
    "hb_he_mulpt_static": (None, [ctx_p, ct_p, ct_p, pt_p]),
    "hb_he_mul": (None, [ctx_p, ct_p, ct_p, ct_p, evk_p]),
    "hb_he_rot": (None, [ctx_p, ct_p, C.c_int, evk_p]),
    "hb_he_conj": (None, [ctx_p, ct_p, evk_p]),
    "hb_he_gemv": (None, [ctx_p, ct_p, vp, ct_p, evk_p]),
    "hb_ctr_hempc_eval": (None, [ctx_p, ct_p, vp, vp, ct_p, ct_p, ct_p, ct_p, evk_p]),
    "hb_ctr_hempc_eval_host": (None, [ctx_p, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, evk_p, vp]),
    "hb_ctr_hempc_submit_host": (C.c_uint64, [ctx_p, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, evk_p, vp]),
    "hb_ctr_hempc_wait": (None, [ctx_p, C.c_uint64]),
    "hb_pk_upload": (None, [ctx_p, evk_p, vp, vp]),
    "hb_sk_upload": (None, [ctx_p, C.POINTER(HbSk), vp]),
    "hb_sk_free": (None, [ctx_p, C.POINTER(HbSk)]),
    "hb_he_enc_pk": (None, [ctx_p, ct_p, vp, C.c_double, evk_p, vp, vp, vp]),
    "hb_he_dec": (None, [ctx_p, vp, ct_p, C.POINTER(HbSk)]),
    "hb_wire_ct_bytes": (C.c_size_t, [ctx_p, C.c_uint, C.c_uint]),
    "hb_wire_evk_bytes": (C.c_size_t, [ctx_p]),
    "hb_ct_serialize": (C.c_size_t, [ctx_p, ct_p, vp, C.c_size_t]),
    "hb_ct_deserialize": (None, [ctx_p, ct_p, vp, C.c_size_t]),
    "hb_evk_serialize": (C.c_size_t, [ctx_p, vp, vp, vp, C.c_size_t]),
    "hb_evk_deserialize": (None, [ctx_p, evk_p, vp, C.c_size_t]),
    "hb_genswk_words": (C.c_uint, [ctx_p]),
    "hb_he_genswk": (None, [ctx_p, C.c_uint, vp, vp, vp, C.POINTER(HbSk), vp, vp]),
    "hb_ntt_forward": (None, [ctx_p, vp, vp, C.c_size_t, C.c_uint, C.c_uint, C.c_int]),
    "hb_ntt_inverse": (None, [ctx_p, vp, vp, C.c_size_t, C.c_uint, C.c_uint]),
    "hb_ntt_forward_host": (None, [ctx_p, vp, vp, C.c_size_t, C.c_uint, C.c_uint, C.c_int]),
    "hb_poly_ntt": (None, [ctx_p, vp, C.c_uint, C.c_size_t, C.c_uint, vp]),
    "hb_rns2words": (None, [ctx_p, vp, C.c_size_t, C.c_uint, C.c_uint, C.c_uint, vp]),
    "hb_poly_rot": (None, [ctx_p, vp, C.c_uint, C.c_size_t, C.c_int, vp]),
}

_LIB = None


def header_symbols() -> list[str]:
    """Every function name ``include/hectr_b200.h`` declares."""
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hb_[a-z0-9_]+)\s*\(", text)))


def load_library() -> C.CDLL:
    """Load ``libhectr_b200.so`` and attach prototypes.  Fails loudly if the extension is not built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(hectr_b200/csrc/build.sh).  There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def exported_symbols() -> list[str]:
    lib = load_library()
    return [n for n in header_symbols() if hasattr(lib, n)]
