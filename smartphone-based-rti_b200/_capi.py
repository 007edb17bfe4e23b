"""ctypes binding of librti_b200.so (the C ABI declared in include/rti_b200.h).

There is no CPU fallback: if the library has not been built (``python
__graft_entry__.py build`` or ``make -C smartphone-based-rti_b200/csrc``) every
call raises ``RtiLibraryMissing``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librti_b200.so")

RTI_OK = 0
RTI_ERR_ARG = -1
RTI_ERR_TOO_FEW = -2
RTI_ERR_CUDA = 1
RTI_ERR_SINGULAR = 2
RTI_ERR_NONFINITE = 3


class RtiLibraryMissing(RuntimeError):
    pass


class RtiError(Exception):
    """Raised for a non-zero status of the C ABI (code in ``.status``)."""

    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status


class RtiSingularError(RtiError, ArithmeticError):
    """>= 1 pixel has a singular system (scipy raises LinAlgError for duplicate RBF nodes)."""


_p = C.c_void_p
_i = C.c_int
_l = C.c_int64
_d = C.c_double
_f = C.c_float

# name -> (restype, argtypes); kept in the order of include/rti_b200.h
SIGNATURES = {
    "rti_version": (_i, []),
    "rti_last_error_string": (C.c_char_p, []),
    "rti_device_info": (_i, [_i, C.POINTER(_i), C.POINTER(_l)]),
    "rti_launch_count": (_l, [_i]),
    "rti_gather_soa": (_i, [_p, _l, _p, _i, _i, _i, _i, _p, _p, _p, _p]),
    "rti_ptm_fit_frames": (_i, [_p, _l, _p, _i, _i, _i, _i, _p, _l, _p, _p]),
    "rti_ptm_fit_frames_masked": (_i, [_p, _l, _p, _i, _i, _i, _i, _i, _i, _p, _l, _p, _p, _p]),
    "rti_ptm_fit_soa": (_i, [_p, _p, _p, _l, _i, _p, _l, _p, _p]),
    "rti_ptm_fit_shared": (_i, [_p, _l, _p, _i, _l, _p, _l, _p]),
    "rti_lights_shared": (_i, [_p, _p, _l, _i, _p, _p]),
    "rti_ptm_shared_pinv": (_i, [_p, _p, _i, _p, _p, _p]),
    "rti_ptm_table": (_i, [_p, _l, _l, _p, _i, _p, _l, _p, _p]),
    "rti_ptm_grids": (_i, [_p, _l, _l, _p, _i, _p, _p]),
    "rti_ptm_fit_table_frames": (_i, [_p, _l, _p, _i, _i, _i, _i, _p, _i, _p, _l, _p, _l, _p, _p]),
    "rti_grids_to_table": (_i, [_p, _l, _i, _p, _l, _p, _p]),
    "rti_rbf_workspace_bytes": (_l, [_i, _i]),
    "rti_rbf_fit_soa": (_i, [_p, _p, _p, _l, _i, _p, _p, _l, _i, _p, _p]),
    "rti_rbf_fit_soa_static": (_i, [_p, _p, _p, _l, _i, _p, _p, _p]),
    "rti_rbf_pcg_workspace_bytes": (_l, [_i, _l]),
    "rti_rbf_fit_pcg": (_i, [_p, _p, _p, _i, _i, _i, _p, _p, _l, _i, _d, _p, _p, _p]),
    "rti_rbf_table": (_i, [_p, _p, _p, _l, _i, _p, _i, _p, _l, _p, _p]),
    "rti_rbf_table_fast": (_i, [_p, _p, _p, _l, _i, _p, _i, _p, _l, _p, _p]),
    "rti_rbf_grids": (_i, [_p, _p, _p, _l, _i, _p, _i, _p, _p]),
    "rti_rbf_shared_operator_elems": (_l, [_i, _i]),
    "rti_rbf_shared_packed_elems": (_l, [_i, _l]),
    "rti_rbf_shared_operator": (_i, [_p, _p, _i, _p, _i, _i, _p, _p, _p, _p, _p]),
    "rti_rbf_shared_pack_frames": (_i, [_p, _l, _i, _l, _p, _p]),
    "rti_rbf_shared_pack_soa": (_i, [_p, _l, _i, _p, _p, _p]),
    "rti_rbf_shared_table": (_i, [_p, _i, _l, _p, _p, _p, _i, _p, _l, _p, _p]),
    "rti_rbf_shared_apply_f64": (_i, [_p, _i, _l, _p, _p, _p, _i, _p, _p]),
    "rti_relight_ptm": (_i, [_p, _l, _l, _i, _d, _d, _p, _p, _p]),
    "rti_relight_ptm_f32": (_i, [_p, _l, _l, _i, _f, _f, _p, _p, _p]),
    "rti_coef_to_f32": (_i, [_p, _l, _p, _p]),
    "rti_relight_from_image": (_i, [_p, _l, _i, _p, _p, _p]),
    "rti_roi_extract": (_i, [_p, _l, _i, _i, _i, _i, _i, _i, _i, _i, _p, _l, _p]),
    "rti_ptm_compute_host": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _i, _p, _p, _i]),
    "rti_table_tile_to_host": (_i, [_p, _l, _l, _l, _p, _l, _l, _p, _p]),
    "rti_table_tile_to_file": (_i, [_p, _l, _l, _l, _i, _p, _l, _p, _p]),
    "rti_host_d2h_bytes": (_l, []),
    "rti_find_magic48": (_l, [_p, _l, C.c_uint64, _p, _l, _i]),
    "rti_copy_bits": (_l, [_p, _l, _l, _p, _l]),
    "rti_destroy": (_i, []),
}

_lib = None


def load():
    """Load the shared library once and declare every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RtiLibraryMissing(
            "librti_b200.so is not built (%s). Build it with `python __graft_entry__.py build`; "
            "this package has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().rti_last_error_string().decode("utf-8", "replace")


def check(status: int, what: str = "") -> None:
    """Turn a C-ABI status into the exception the reference's callers would see."""
    if status == RTI_OK:
        return
    msg = last_error() or what
    if status == RTI_ERR_TOO_FEW:
        # the reference fails in np.dot with a shape mismatch (analysis.py:296-298)
        raise ValueError(msg)
    if status == RTI_ERR_SINGULAR:
        raise RtiSingularError(status, msg)
    if status == RTI_ERR_NONFINITE:
        # analysis.py:407: ValueError (NaN/inf) or OverflowError (out of int32)
        raise ValueError(msg)
    raise RtiError(status, msg)


def call(name: str, *args) -> None:
    # numpy integers (e.g. an index taken from an array) are not accepted by ctypes
    args = [int(a) if hasattr(a, "__index__") and not isinstance(a, (int, bool)) else a for a in args]
    check(getattr(load(), name)(*args), name)
