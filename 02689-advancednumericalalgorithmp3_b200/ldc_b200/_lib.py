"""ctypes binding of the C ABI in include/ldc_b200.h.

The shared library is built in-tree by `__graft_entry__.build()` (or `python -m ldc_b200.build`).
There is no fallback: if the library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libldc_b200.so")

LDC_PATH_AUTO, LDC_PATH_TILED, LDC_PATH_RESIDENT, LDC_PATH_RESIDENT_EO, LDC_PATH_STREAM_EO = 0, 1, 2, 3, 4
LDC_PATH_RESIDENT_EO_SPLIT = 5
LDC_PATH_STREAM_TMA = 6

_dp = C.POINTER(C.c_double)
_vp = C.c_void_p

# name -> (restype, argtypes); mirrors include/ldc_b200.h one to one
SIGNATURES = {
    "ldc_version": (C.c_char_p, []),
    "ldc_last_error": (C.c_char_p, []),
    "ldc_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, C.c_int, C.c_int, _vp]),
    "ldc_destroy": (None, [_vp]),
    "ldc_set_path": (C.c_int, [_vp, C.c_int]),
    "ldc_resident_supported": (C.c_int, [_vp]),
    "ldc_kernel_name": (C.c_char_p, [_vp]),
    "ldc_set_operators": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_double, C.c_double]),
    "ldc_set_cases": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ldc_set_state": (C.c_int, [_vp, _vp, _vp, _vp]),
    "ldc_get_state": (C.c_int, [_vp, _vp, _vp, _vp]),
    "ldc_init_state": (C.c_int, [_vp]),
    "ldc_residual": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "ldc_step": (C.c_int, [_vp, C.c_int64, C.c_int, _vp]),
    "ldc_solve": (C.c_int, [_vp, _vp, C.c_int64, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp]),
    "ldc_solve_segment": (C.c_int, [_vp, _vp, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp,
                                    _vp]),
    "ldc_solve_history": (C.c_int, [_vp, _vp, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, _vp, _vp, C.c_int64, _vp, _vp,
                                    _vp, _vp]),
    "ldc_solve_sampled": (C.c_int, [_vp, _vp, C.c_int64, C.c_int, C.c_int, C.c_int, _vp, C.c_int64, _vp, _vp, _vp, _vp, _vp,
                                    _vp]),
    "ldc_diagnostics": (C.c_int, [_vp, _vp, _vp]),
    "ldc_vortex_metrics": (C.c_int, [_vp] + [_vp] * 9),
    "ldc_validation_errors": (C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ldc_prolongate": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int]),
    "ldc_restrict": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "ldc_launch_count": (C.c_int64, [_vp]),
    "ldc_last_kernel_ms": (C.c_double, [_vp]),
    "ldc_fp64_peak": (C.c_int, [C.c_int, C.c_int, _dp]),
    "ldc_diagnostics_noalloc": (C.c_int, [C.c_int, _vp, C.c_int, C.c_int, C.c_int] + [_vp] * 6),
}

_lib = None


class LdcError(RuntimeError):
    """A call into libldc_b200 returned a non-zero status."""


def load():
    """Load libldc_b200.so (once) and declare every prototype.  Raises if the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LdcError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). ldc_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().ldc_last_error().decode("utf-8", "replace")
        raise LdcError(f"{what or 'ldc call'} failed (code {rc}): {msg}")


def ptr(a):
    """Raw address of a NumPy array / torch tensor / None, as c_void_p."""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):  # torch tensor (device or host)
        if not a.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return C.c_void_p(a.data_ptr())
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError("array must be C-contiguous")
    return C.c_void_p(a.ctypes.data)
