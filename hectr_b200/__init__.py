"""hectr_b200 -- B200 (sm_100a) CKKS evaluation engine behind the GPQHE/HECTR C API.

The product is the C-ABI shared library ``libhectr_b200.so`` (``include/hectr_b200.h``); this
package is the thin ctypes host-side mirror used by the tests, ``bench.py`` and
``__graft_entry__``.  There is no CPU fallback: importing works anywhere (so the symbol
table can be checked), but creating an :class:`Engine` without a B200 aborts loudly.
"""
from ._lib import LIB_PATH, load_library, exported_symbols  # noqa: F401
from .engine import Engine, Ciphertext, EvalKeys  # noqa: F401

__all__ = ["Engine", "Ciphertext", "EvalKeys", "load_library", "exported_symbols", "LIB_PATH"]
