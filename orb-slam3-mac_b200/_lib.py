"""ctypes binding of liborbx_b200.so (the C ABI of include/orbx.h and include/orbm.h).

The product path is the CUDA library: if it is missing this module raises -- there is no CPU
fallback and no routing through oracle/.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# ORBX_B200_LIB: developer override used for A/B runs of kernel variants (same ABI)
LIB_PATH = os.environ.get("ORBX_B200_LIB") or os.path.join(HERE, "liborbx_b200.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])

ORBX_OK = 0
ORBX_E_EMPTY_IMAGE = -1
ORBX_E_BAD_ARG = -2
ORBX_E_BAD_TYPE = -3
ORBX_E_CAPACITY = -4
ORBX_E_TOO_LARGE = -5
ORBX_E_UNSUPPORTED = -6
ORBX_E_CUDA = -7
ORBX_E_BUSY = -8


class OrbxConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("max_width", C.c_int32),
                ("max_height", C.c_int32), ("max_batch", C.c_int32), ("n_slots", C.c_int32),
                ("device", C.c_int32), ("flags", C.c_uint32)]


class OrbxPyramidView(C.Structure):
    """orbx_pyramid_view (include/orbx.h): device pointers of one frame's pyramid."""
    _fields_ = [("data", C.c_void_p * 16), ("pitch", C.c_int32 * 16), ("width", C.c_int32 * 16),
                ("height", C.c_int32 * 16), ("nlevels", C.c_int32), ("device", C.c_int32)]


class OrbxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"orbx status {code}: {msg}")
        self.code = code


_lib = None

# every symbol include/orbx.h declares: (name, restype, argtypes)
_vp, _i, _sz, _ip = C.c_void_p, C.c_int, C.c_size_t, C.c_void_p
ORBX_SYMBOLS = {
    "orbx_create": (_i, [C.POINTER(OrbxConfig), C.POINTER(_vp)]),
    "orbx_destroy": (_i, [_vp]),
    "orbx_last_error_string": (C.c_char_p, []),
    "orbx_abi_version": (_i, []),
    "orbx_get_tables": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_max_keypoints": (_i, [_vp]),
    "orbx_device_pyramid": (_i, [_vp, _i, _i, C.POINTER(OrbxPyramidView)]),
    "orbx_extract": (_i, [_vp, _vp, _i, _i, _sz, _i, _i, _vp, _vp, _i, _ip, _ip]),
    "orbx_extract_batch": (_i, [_vp, _vp, _i, _i, _i, _sz, _vp, _vp, _vp, _i, _ip, _ip, _ip]),
    "orbx_submit_batch": (_i, [_vp, _i, _vp, _i, _i, _i, _sz, _vp, _vp, _vp, _i, _ip, _ip, _ip]),
    "orbx_wait": (_i, [_vp, _i]),
    "orbx_host_alloc": (_vp, [_sz]),
    "orbx_host_free": (None, [_vp]),
    "orbx_extract_device": (_i, [_vp, _i, _vp, _i, _i, _i, _sz, _sz, _vp, C.POINTER(_vp),
                                 C.POINTER(_vp), C.POINTER(_vp)]),
    "orbx_slot_stream": (_vp, [_vp, _i]),
    "orbx_set_input_format": (_i, [_vp, _i, _i]),
    "orbx_set_rectify_maps": (_i, [_vp, _vp, _vp, _i, _i, C.c_size_t]),
    "orbx_level_size": (_i, [_vp, _i, _i, _i, _ip, _ip]),
    "orbx_get_pyramid_level": (_i, [_vp, _i, _i, _i, _i, _vp, _sz]),
    "orbx_debug_candidates": (_i, [_vp, _i, _i, _i, _vp, _i]),
    "orbx_debug_survivors": (_i, [_vp, _i, _i, _i, _vp, _i]),
    "orbx_last_launch_count": (_i, [_vp, _i]),
    "orbx_last_stage_ms": (_i, [_vp, _i, _vp, _i]),
    "orbx_sink_create": (_i, [_i, _i, _i, C.POINTER(_vp)]),
    "orbx_sink_export": (_i, [_vp, _vp]),
    "orbx_sink_open": (_i, [_i, _vp, _i, _i, C.POINTER(_vp)]),
    "orbx_sink_destroy": (_i, [_vp]),
    "orbx_sink_pointers": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)]),
    "orbx_sink_read_frame": (_i, [_vp, _i, _vp, _vp, _vp, _i]),
    "orbx_set_result_sink": (_i, [_vp, _i, _vp, _i]),
    "orbx_debug_sincos_check": (_i, [_i, C.c_float, C.c_float, C.POINTER(C.c_ulonglong), C.POINTER(C.c_float)]),
    "orbx_set_result_sink_ex": (_i, [_vp, _i, _vp, _i, _i]),
    "orbx_sink_wait": (_i, [_vp, _i, C.c_uint32, _vp, _i]),
    "orbx_sink_release": (_i, [_vp, _i, C.c_uint32, _vp]),
    "orbx_sink_epoch": (_i, [_vp, _i, C.POINTER(C.c_uint32), C.POINTER(_i)]),
}


def load():
    """Loads the CUDA library; raises if it has not been built (python __graft_entry__.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                          f"g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in ORBX_SYMBOLS.items():
        f = getattr(lib, name)
        f.restype = res
        f.argtypes = args
    try:
        from . import _lib_match
        _lib_match.declare(lib)
    except ImportError:
        pass
    _lib = lib
    return lib


def last_error() -> str:
    return load().orbx_last_error_string().decode("utf-8", "replace")


def check(code: int):
    if code != ORBX_OK:
        raise OrbxError(code, last_error())
