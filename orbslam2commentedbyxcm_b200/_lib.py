"""ctypes binding of liborbx.so (include/orbx.h).  There is no CPU fallback: if the library is missing or no CUDA
device is usable, every compute call raises."""
import ctypes as C
import os

import numpy as np

from . import _build

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")]
)

NUM_STAGES = 7
STAGE_NAMES = ("h2d", "pyramid", "fast", "octree", "blur", "desc", "d2h")


class OrbxError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int),
                ("ini_th_fast", C.c_int), ("min_th_fast", C.c_int)]


class FrameView(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.c_void_p), ("y", C.c_void_p), ("octave", C.c_void_p), ("u_right", C.c_void_p),
                ("desc", C.c_void_p), ("occupied", C.c_void_p), ("cell_start", C.c_void_p), ("cell_items", C.c_void_p),
                ("min_x", C.c_float), ("min_y", C.c_float), ("grid_w_inv", C.c_float), ("grid_h_inv", C.c_float),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int)]


class MapPointView(C.Structure):
    _fields_ = [("n", C.c_int), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p),
                ("level", C.c_void_p), ("view_cos", C.c_void_p), ("valid", C.c_void_p), ("obs_pos", C.c_void_p),
                ("desc", C.c_void_p)]


class WindowQueryView(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.c_void_p), ("y", C.c_void_p), ("radius", C.c_void_p), ("min_level", C.c_void_p),
                ("max_level", C.c_void_p), ("xr", C.c_void_p), ("angle", C.c_void_p), ("valid", C.c_void_p),
                ("claims_block", C.c_void_p), ("desc", C.c_void_p)]


class FeatVecView(C.Structure):
    _fields_ = [("n_nodes", C.c_int), ("node_id", C.c_void_p), ("node_start", C.c_void_p), ("node_feat", C.c_void_p)]


class KFView(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.c_void_p), ("y", C.c_void_p), ("octave", C.c_void_p), ("angle", C.c_void_p),
                ("u_right", C.c_void_p), ("desc", C.c_void_p), ("has_mp", C.c_void_p), ("fv", C.POINTER(FeatVecView))]


_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# every symbol include/orbx.h declares: (restype, argtypes)
SIGNATURES = {
    "orbx_last_error": (C.c_char_p, []),
    "orbx_device_count": (_i, []),
    "orbx_create": (_i, [C.POINTER(Config), _i, C.POINTER(_vp)]),
    "orbx_destroy": (None, [_vp]),
    "orbx_get_tables": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_max_keypoints": (_i, [_vp, _i, _i]),
    "orbx_extract_batch": (_i, [_vp, _vp, _i, _i, _i, _sz, _sz, _vp, _vp, _i, _vp]),
    "orbx_alloc_pinned": (_vp, [_sz]),
    "orbx_free_pinned": (None, [_vp]),
    "orbx_extract": (_i, [_vp, _vp, _i, _i, _sz, _vp, _vp, _i, _vp]),
    "orbx_extract_batch_device": (_i, [_vp, _vp, _i, _i, _i, _sz, _sz, _vp, _vp, _i, _vp]),
    "orbx_sync": (_i, [_vp]),
    "orbx_stream": (_vp, [_vp]),
    "orbx_get_level_size": (_i, [_vp, _i, _vp, _vp]),
    "orbx_get_pyramid_level": (_i, [_vp, _i, _i, _vp, _sz]),
    "orbx_get_pyramid_all": (_i, [_vp, _i, _vp, _vp, _vp]),
    "orbx_set_profiling": (_i, [_vp, _i]),
    "orbx_stage_times": (_i, [_vp, _vp]),
    "orbx_last_launch_count": (_i, [_vp]),
    "orbx_set_use_graph": (_i, [_vp, _i]),
    "orbx_debug_get_candidates": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "orbx_debug_get_blurred": (_i, [_vp, _i, _i, _vp]),
    "orbx_debug_sincosf_checksum": (_i, [_i, C.c_uint32, C.c_uint64, C.c_uint64, _vp, _vp]),
    "orbx_debug_sincosf": (_i, [_i, C.c_uint32, _i, _vp, _vp]),
    "orbx_debug_fast_atan2_checksum": (_i, [_i, _i, C.c_uint64, C.c_uint64, C.c_uint64, _vp]),
    "orbx_debug_fast_atan2": (_i, [_i, _i, C.c_uint64, _i, _vp]),
    "orbm_descriptor_distance": (_i, [_vp, _vp]),
    "orbm_create": (_i, [_i, C.POINTER(_vp)]),
    "orbm_destroy": (None, [_vp]),
    "orbm_stream": (_vp, [_vp]),
    "orbm_sync": (_i, [_vp]),
    "orbm_bruteforce": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp, _i]),
    "orbm_bruteforce_noalloc": (_i, [_i, _vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "orbm_db_create": (_i, [_vp, _vp, _i, _i, C.POINTER(_vp)]),
    "orbm_db_destroy": (None, [_vp]),
    "orbm_bruteforce_db": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i]),
    "orbm_bruteforce_sharded_db": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "orbm_merge_partials": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "orbm_undistort_keypoints": (_i, [_vp, _vp, _vp, _i, _f, _f, _f, _f, _vp, _i, _vp, _vp, _i]),
    "orbm_assign_grid": (_i, [_vp, _vp, _vp, _i, _f, _f, _f, _f, _vp, _vp, _i]),
    "orbm_exchange_create": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "orbm_exchange_connect": (_i, [_vp, _vp]),
    "orbm_exchange_connect_local": (_i, [_vp, _vp]),
    "orbm_exchange_destroy": (None, [_vp]),
    "orbm_bruteforce_sharded": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "orbm_bruteforce_sharded_phases": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp, _i]),
    "orbm_exchange_timed_out": (_i, [_vp]),
    "orbm_debug_exchange_snapshot": (_i, [_vp, _vp, C.c_size_t, _vp]),
    "orbm_search_by_projection": (_i, [_vp, C.POINTER(FrameView), C.POINTER(MapPointView), _f, _f, _vp, _vp]),
    "orbm_search_window": (_i, [_vp, C.POINTER(FrameView), _vp, C.POINTER(WindowQueryView), _i, _i, _f, _i, _vp, _vp]),
    "orbm_search_best": (_i, [_vp, C.POINTER(FrameView), C.POINTER(WindowQueryView), _i, _vp, _i, _vp, _vp]),
    "orbm_search_by_sim3": (_i, [_vp, C.POINTER(FrameView), C.POINTER(WindowQueryView), C.POINTER(FrameView),
                                 C.POINTER(WindowQueryView), _vp, _vp]),
    "orbm_search_for_initialization": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, C.POINTER(FrameView), _vp, _i, _f, _i, _vp, _vp]),
    "orbm_search_for_triangulation": (_i, [_vp, C.POINTER(KFView), C.POINTER(KFView), _vp, _f, _f, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "orbm_stereo_rows": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _f, _f, _vp, _vp]),
    "orbm_stereo_match": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp]),
    "orbm_set_stereo_outlier_mode": (_i, [_vp, _i]),
    "orbm_search_by_bow_kf": (_i, [_vp, _i, _vp, _vp, _vp, C.POINTER(FeatVecView), _i, _vp, _vp, _vp, C.POINTER(FeatVecView), _f, _i, _vp, _vp]),
    "orbm_vocab_create": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, C.POINTER(_vp)]),
    "orbm_vocab_destroy": (None, [_vp]),
    "orbm_bow_transform": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "orbm_distinctive_descriptors": (_i, [_vp, _i, _vp, _vp, _vp]),
    "orbm_search_by_bow": (_i, [_vp, _i, _vp, _vp, _vp, C.POINTER(FeatVecView), _i, _vp, _vp, C.POINTER(FeatVecView), _f, _i, _vp, _vp]),
}

_lib = None


def lib_path():
    return _build.LIB


def load():
    """Load liborbx.so (never builds implicitly: a missing library is an error, not a fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise OrbxError(
            "liborbx.so is missing (%s). Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `python -m orbslam2commentedbyxcm_b200._build`; there is no CPU fallback." % path
        )
    L = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError if the C-ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        msg = load().orbx_last_error().decode("utf-8", "replace")
        raise OrbxError("%s failed with code %d: %s" % (what, rc, msg))
