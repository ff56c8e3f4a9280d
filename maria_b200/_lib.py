"""ctypes binding of libmaria_atm.so (include/maria_atm.h).  There is no CPU fallback: if the library is missing
or a call fails, the caller gets an exception."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _build

OK, ERR_CUDA, ERR_ARG, ERR_OUT_OF_GRID, ERR_TABLE_RANGE = 0, 1, 2, 3, 4
F32, F64 = 0, 1
UNITS_PW, UNITS_KRJ = 0, 1
MAX_LAYERS, MAX_BANDS = 64, 16

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)
vp = C.c_void_p


class LayerDesc(C.Structure):
    _fields_ = [
        ("data_dev", vp), ("n_ext", C.c_int32), ("n_cs", C.c_int32), ("h", C.c_double), ("pwv_rms", C.c_double),
        ("transform", C.c_double * 9), ("ext0", C.c_double), ("ext_step", C.c_double), ("cs0", C.c_double),
        ("cs_last", C.c_double), ("vx_host", c_dp), ("vy_host", c_dp),
    ]


# name -> (restype, argtypes); every symbol include/maria_atm.h declares
SIGNATURES = {
    "matm_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(vp)]),
    "matm_destroy": (C.c_int, [vp]),
    "matm_last_error": (C.c_int, [vp, C.c_char_p, C.c_int]),
    "matm_abi_version": (C.c_int, []),
    "matm_fill_normal": (C.c_int, [vp, vp, C.c_size_t, C.c_uint64, C.c_uint64, vp]),
    "matm_ar_generate": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, vp]),
    "matm_smooth_layer": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, vp, vp, vp]),
    "matm_set_scan": (C.c_int, [vp, C.c_int, c_dp, C.c_double, c_dp, c_dp, C.c_int64, c_dp, c_dp, c_dp]),
    "matm_set_layers": (C.c_int, [vp, C.c_int, C.POINTER(LayerDesc), C.c_double]),
    "matm_set_layer_data": (C.c_int, [vp, C.c_int, C.POINTER(vp)]),
    "matm_set_detectors": (C.c_int, [vp, C.c_int64, c_dp, c_dp, c_ip, c_dp, c_dp]),
    "matm_set_gain": (C.c_int, [vp, c_dp, vp]),
    "matm_set_bands": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp,
                                 c_dp, C.c_double, C.c_double, C.c_double, c_dp]),
    "matm_zenith_pwv": (C.c_int, [vp, C.c_int64, C.c_int64, vp, vp]),
    "matm_upsample_linear": (C.c_int, [vp, vp, C.c_int64, vp, vp]),
    "matm_synthesize": (C.c_int, [vp, C.c_int64, C.c_int64, vp, C.c_int, C.c_int, vp]),
    "matm_set_tuning": (C.c_int, [vp, C.c_int, C.c_int]),
    "matm_set_timing": (C.c_int, [vp, C.c_int]),
    "matm_get_timing": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "matm_synthesize_host": (C.c_int, [vp, C.c_int64, C.c_int64, vp, C.c_int, C.c_int]),
    "matm_synthesize_noalloc": (C.c_int, [vp, C.c_int64, C.c_int64, vp, C.c_int, C.c_int, vp]),
    "matm_calibrate": (C.c_int, [vp, C.c_int64, C.c_int64, vp, vp, C.c_int, vp]),
    "matm_transmission": (C.c_int, [vp, C.c_int64, C.c_int64, C.c_double, C.c_double, vp, vp]),
    "matm_noise_scale": (C.c_int, [vp, C.c_int64, C.c_int64, vp, vp, c_dp, c_dp, vp, vp]),
    "matm_pointing": (C.c_int, [vp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, vp, vp, vp]),
    "matm_check_flags": (C.c_int, [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64),
                                   C.POINTER(C.c_int64)]),
    "matm_launch_count": (C.c_int64, [vp]),
}

_lib = None


def library_path():
    return _build.LIB


def load(build_if_missing=True):
    """dlopen libmaria_atm.so and bind every symbol.  Raises if the library cannot be found or built."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise RuntimeError(f"{path} is missing: run `python -m maria_b200._build`")
        _build.build()
    elif _build.is_stale():
        import logging

        logging.getLogger("maria_b200").warning(f"{path} is older than its sources: run `python -m maria_b200._build` "
                                                "(measurements would describe the old kernels)")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def dptr(a):
    """float64 host array -> POINTER(c_double) (the array must stay alive during the call)."""
    if a is None:
        return None
    assert isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(c_dp)


def iptr(a):
    assert isinstance(a, np.ndarray) and a.dtype == np.int32 and a.flags.c_contiguous
    return a.ctypes.data_as(c_ip)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)
