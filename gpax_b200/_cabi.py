"""ctypes binding of ``include/gpax_b200.h`` (the C-ABI drop-in boundary).

The shared library is built in-tree by ``gpax_b200/csrc/build.sh`` (or
``__graft_entry__.build()``).  There is no fallback: if the library is missing
or a symbol is absent, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libgpax_b200.so")

RBF, MATERN52, PERIODIC = 0, 1, 2
ACQ_EI, ACQ_UCB, ACQ_UE, ACQ_MEAN = 0, 1, 2, 3
FULL, LOWER = 0, 1
NB = 128
STATIC_SCRATCH_BYTES = 8 << 20      # GPX_STATIC_SCRATCH_BYTES (gpx_predict_static)
MAX_D = 8
KIND_BY_NAME = {"rbf": RBF, "matern52": MATERN52, "periodic": PERIODIC}


class KernelParams(C.Structure):
    """``gpx_kernel_params``."""
    _fields_ = [
        ("kind", C.c_int32), ("d", C.c_int32), ("dl", C.c_int32), ("dp", C.c_int32),
        ("scale", C.c_void_p), ("length", C.c_void_p), ("period", C.c_void_p),
    ]


class GpxError(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with gpax_b200/csrc/build.sh "
        "(python -c 'import __graft_entry__ as g; g.build()'). gpax_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH)

_i64, _i32, _dbl, _ptr, _sz = C.c_int64, C.c_int32, C.c_double, C.c_void_p, C.c_size_t
_kp = C.POINTER(KernelParams)

# name -> (restype, argtypes); must list every symbol include/gpax_b200.h declares
SIGNATURES = {
    "gpx_abi_version": (C.c_int, []),
    "gpx_last_error_string": (C.c_char_p, []),
    "gpx_launch_count": (C.c_uint64, []),
    "gpx_gram": (C.c_int, [_kp, _i64, _i64, _i64, _ptr, _ptr, _ptr, _i64, _i64, _ptr, _i32,
                           _i32, _ptr, _i64, _ptr]),
    "gpx_dinv_bytes": (_sz, [_i64, _i64]),
    "gpx_potrf_schedule": (_i32, [_i64, _i64]),
    "gpx_potrf_batched": (C.c_int, [_i64, _i64, _ptr, _i64, _ptr, _ptr, _ptr]),
    "gpx_trtri_batched": (C.c_int, [_i64, _i64, _ptr, _i64, _ptr, _ptr, _i64, _ptr]),
    "gpx_mll_ws_bytes": (_sz, [_i64, _i64]),
    "gpx_mll": (C.c_int, [_i64, _i64, _ptr, _i64, _ptr, _ptr, _i64, _ptr, _ptr, _ptr, _ptr]),
    "gpx_mll_grad_ws_bytes": (_sz, [_i64, _i64, _i32]),
    "gpx_mll_grad": (C.c_int, [_kp, _i64, _i64, _ptr, _ptr, _i64, _ptr, _ptr, _ptr, _ptr, _ptr,
                               _ptr, _ptr]),
    "gpx_predict_ws_bytes": (_sz, [_i64, _i64, _i32]),
    "gpx_predict": (C.c_int, [_kp, _i64, _i64, _i64, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _dbl,
                              _ptr, _ptr, _ptr, _ptr, _ptr]),
    "gpx_predict_i8_ws_bytes": (_sz, [_i64, _i64, _i32]),
    "gpx_predict_i8": (C.c_int, [_kp, _i64, _i64, _i64, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _dbl,
                                 _ptr, _ptr, _ptr, _i32, _ptr, _ptr]),
    "gpx_predict_i8_want": (_sz, [_kp, _i64, _i64, _i64]),
    "gpx_predict_i8_attach": (None, [_ptr, _sz, _i32]),
    "gpx_predict_static": (C.c_int, [_kp, _i64, _i64, _i64, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _dbl,
                                     _ptr, _ptr, _ptr, _ptr]),
    "gpx_acq_parts": (_i64, [_i64]),
    "gpx_acq_reduce": (C.c_int, [_i64, _i64, _ptr, _ptr, _i32, _dbl, _dbl, _dbl, _dbl, _dbl,
                                 _i64, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "gpx_argmax_finalize": (C.c_int, [_i64, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "gpx_predict_acq_tile": (_i64, [_i64, _i64]),
    "gpx_predict_acq_ws_bytes": (_sz, [_i64, _i64, _i32, _i64]),
    "gpx_predict_acq": (C.c_int, [_kp, _i64, _i64, _i64, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _dbl,
                                  _ptr, _i32, _dbl, _dbl, _dbl, _dbl, _dbl, _i64, _i64, _ptr, _ptr,
                                  _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "gpx_halton": (C.c_int, [_i64, _i32, _i64, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "gpx_threefry2x32": (C.c_int, [_i64, C.c_uint32, C.c_uint32, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "gpx_random_split": (C.c_int, [C.c_uint32, C.c_uint32, _i64, _i32, _ptr, _ptr]),
    "gpx_random_normal": (C.c_int, [_ptr, _i64, _i64, _i64, _i32, _ptr, _i64, _ptr]),
    "gpx_draw_ws_bytes": (_sz, [_i64, _i64, _i64, _i64]),
    "gpx_draw": (C.c_int, [_kp, _i64, _i64, _i64, _i64, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _ptr,
                           C.c_uint32, C.c_uint32, _i32, _ptr, _ptr, _ptr, _ptr]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)          # AttributeError if the symbol is missing
    _fn.restype = _res
    _fn.argtypes = _args

ABI_VERSION = 10
if lib.gpx_abi_version() != ABI_VERSION:
    raise ImportError("libgpax_b200.so ABI version mismatch")


def check(rc: int, what: str = "") -> None:
    """Raise on a non-zero C-ABI return code (negative: bad argument; positive:
    cudaError_t)."""
    if rc != 0:
        msg = lib.gpx_last_error_string().decode("utf-8", "replace")
        raise GpxError(f"{what} failed with code {rc}: {msg}")
