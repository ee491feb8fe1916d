"""ctypes prototypes and POD structs of include/orbm.h."""
from __future__ import annotations

import ctypes as C

import numpy as np


class FrameView(C.Structure):  # orbm_frame_view
    _fields_ = [("keys", C.c_void_p), ("desc", C.c_void_p), ("n", C.c_int32),
                ("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("grid_w_inv", C.c_float), ("grid_h_inv", C.c_float),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int32), ("u_right", C.c_void_p)]


MAP_POINT_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_xr", "<f4"), ("view_cos", "<f4"),
                            ("depth", "<f4"), ("level", "<i4"), ("observations", "<i4"),
                            ("in_view", "u1"), ("bad", "u1"), ("_pad", "u1", (2,))])
LAST_OBS_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("inv_z", "<f4"), ("angle", "<f4"),
                           ("octave", "<i4"), ("observations", "<i4"), ("valid", "u1"), ("_pad", "u1", (3,))])
_fv = C.POINTER(FrameView)
CAMERA_DTYPE = np.dtype([("fx", "<f4"), ("fy", "<f4"), ("cx", "<f4"), ("cy", "<f4"), ("k1", "<f4"), ("k2", "<f4"),
                         ("p1", "<f4"), ("p2", "<f4"), ("k3", "<f4")])  # orbm_camera


class BowView(C.Structure):  # orbm_bow_view
    _fields_ = [("keys", C.c_void_p), ("desc", C.c_void_p), ("n", C.c_int32), ("nnodes", C.c_int32),
                ("node_ids", C.c_void_p), ("node_ptr", C.c_void_p), ("node_items", C.c_void_p)]


class FrameRight(C.Structure):  # orbm_frame_right
    _fields_ = [("keys", C.c_void_p), ("desc", C.c_void_p), ("n", C.c_int32),
                ("left_to_right", C.c_void_p), ("right_to_left", C.c_void_p)]


MAP_POINT_RIGHT_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("view_cos", "<f4"), ("level", "<i4"),
                                  ("in_view", "u1"), ("_pad", "u1", (3,))])  # orbm_map_point_right
assert MAP_POINT_RIGHT_DTYPE.itemsize == 20
AREA_QUERY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"),
                             ("valid", "<i4"), ("desc_row", "<i4")])  # orbm_area_query
assert MAP_POINT_DTYPE.itemsize == 32 and LAST_OBS_DTYPE.itemsize == 28 and AREA_QUERY_DTYPE.itemsize == 28

PROJ_POINT_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("ur", "<f4"), ("angle", "<f4"), ("level", "<i4"),
                             ("valid", "u1"), ("_pad", "u1", (3,))])  # orbm_proj_point
assert PROJ_POINT_DTYPE.itemsize == 24
PAIR_TEST = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.c_int32)  # orbm_pair_test

_vp, _i, _f = C.c_void_p, C.c_int, C.c_float
# limits of orbm_area_candidates_static (include/orbm.h)
ORBM_STATIC_MAX_KEYS = 16384
ORBM_STATIC_MAX_QUERIES = 32768

ORBM_SYMBOLS = {
    "orbm_create": (_i, [_i, C.POINTER(_vp)]),
    "orbm_destroy": (_i, [_vp]),
    "orbm_hamming_csr": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp, _i, _vp]),
    "orbm_knn2": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp]),
    "orbm_stereo_rows": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _f, _f, _vp, _vp]),
    "orbm_compute_stereo_matches": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _f, _f, _vp, _vp]),
    "orbm_search_by_projection_points": (_i, [_vp, _fv, _vp, _vp, _i, _f, _i, _f, _f, _vp, _vp]),
    "orbm_search_by_projection_frames": (_i, [_vp, _fv, _vp, _vp, _i, _f, _i, _i, _f, _i, _vp, _vp]),
    "orbm_search_by_projection_points_stereo": (_i, [_vp, _fv, C.POINTER(FrameRight), _vp, _vp, _vp, _i, _f, _i, _f, _f,
                                                     _vp, _vp]),
    "orbm_search_by_projection_frames_stereo": (_i, [_vp, _fv, C.POINTER(FrameRight), _vp, _vp, _vp, _i, _f, _i, _i, _i,
                                                     _vp, _vp]),
    "orbm_search_by_bow": (_i, [_vp, C.POINTER(BowView), _vp, C.POINTER(BowView), _i, _f, _i, _vp]),
    "orbm_search_for_initialization": (_i, [_vp, _fv, _fv, _vp, _vp, _i, _f, _i]),
    "orbm_search_by_bow_kf": (_i, [_vp, C.POINTER(BowView), _vp, _i, C.POINTER(BowView), _vp, _i, _f, _i, _vp]),
    "orbm_search_by_projection_sim3": (_i, [_vp, _fv, _vp, _vp, _i, _i, _f, _vp, _vp]),
    "orbm_search_by_projection_reloc": (_i, [_vp, _fv, _vp, _vp, _i, _f, _i, _i, _vp, _vp]),
    "orbm_fuse": (_i, [_vp, _fv, _vp, _vp, _i, _f, _vp, _i, _vp, _vp]),
    "orbm_search_by_sim3": (_i, [_vp, _fv, _fv, _vp, _vp, _vp, _vp, _f, _vp]),
    "orbm_search_for_triangulation": (_i, [_vp, C.POINTER(BowView), _vp, _vp, C.POINTER(BowView), _vp, _vp, _i, _vp,
                                           _vp, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    "orbm_distinctive_descriptors": (_i, [_vp, _vp, _vp, _i, _vp]),
    "orbm_vocab_create": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, C.POINTER(_vp)]),
    "orbm_vocab_destroy": (_i, [_vp]),
    "orbm_bow_transform": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, C.POINTER(C.c_int32), _vp, _vp, _vp,
                                C.POINTER(C.c_int32)]),
    "orbm_area_candidates": (_i, [_vp, _fv, _vp, _i, _vp, _i, _vp, _vp, _vp, _i]),
    "orbm_frame_upload": (_i, [_vp, _fv, C.POINTER(_vp)]),
    "orbm_frame_from_device": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _vp, _i, C.POINTER(_vp)]),
    "orbm_frame_get_view": (_vp, [_vp]),
    "orbm_frame_set_u_right": (_i, [_vp, _vp]),
    "orbm_frame_destroy": (_i, [_vp]),
    "orbm_bank_create": (_i, [_vp, _i, C.POINTER(C.c_void_p)]),
    "orbm_bank_write": (_i, [_vp, _i, _i, _vp]),
    "orbm_bank_rows": (_i, [_vp]),
    "orbm_bank_destroy": (_i, [_vp]),
    "orbm_frame_bank": (_vp, [_vp]),
    "orbm_search_by_projection_points_bank": (_i, [_vp, _fv, _vp, _vp, _vp, _i, _f, _i, _f, _f, _vp, _vp]),
    "orbm_search_by_projection_frames_bank": (_i, [_vp, _fv, _vp, _vp, _vp, _i, _f, _i, _i, _f, _i, _vp, _vp]),
    "orbm_area_candidates_static": (_i, [_i, _vp, _vp, _vp, _i, _f, _f, _f, _f, _vp, _i, _vp, _vp, _vp, _vp, _i]),
    "orbm_last_device_ms": (_i, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_longlong)]),
    "orbm_measure_popc_peak": (_i, [_vp, C.POINTER(C.c_double)]),
}


def declare(lib):
    for name, (res, args) in ORBM_SYMBOLS.items():
        f = getattr(lib, name)
        f.restype = res
        f.argtypes = args
