"""ctypes binding of liborbx.so (include/orbx.h).  Loads the in-tree CUDA library and fails loudly
if it is missing: there is no CPU fallback anywhere in this package."""
import ctypes as C
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "liborbx.so"

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

GRID_COLS, GRID_ROWS = 64, 48
GRID_CELLS = GRID_COLS * GRID_ROWS

OK, ERR_ARG, ERR_EMPTY_IMAGE, ERR_CTOR, ERR_CUDA, ERR_CAPACITY, ERR_ASPECT, ERR_STATE = 0, -1, -2, -3, -4, -5, -6, -7
_NAMES = {ERR_ARG: "ORBX_ERR_ARG", ERR_EMPTY_IMAGE: "ORBX_ERR_EMPTY_IMAGE", ERR_CTOR: "ORBX_ERR_CTOR",
          ERR_CUDA: "ORBX_ERR_CUDA", ERR_CAPACITY: "ORBX_ERR_CAPACITY", ERR_ASPECT: "ORBX_ERR_ASPECT",
          ERR_STATE: "ORBX_ERR_STATE"}


class OrbxError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__("%s (%d) %s" % (_NAMES.get(code, "?"), code, msg))
        self.code = code


class Frame(C.Structure):  # orbx_frame
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("desc", C.c_void_p), ("grid_offsets", C.c_void_p),
                ("grid_indices", C.c_void_p)]


vp, ci, cf, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
# name -> (restype, argtypes): every symbol include/orbx.h declares
SIGNATURES = {
    "orbx_create": (vp, [ci, ci, ci, ci, cf, ci, ci, ci, ci, vp]),
    "orbx_destroy": (None, [vp]),
    "orbx_last_error": (C.c_char_p, [vp]),
    "orbx_set_trig_mode": (ci, [vp, ci]),
    "orbx_sincosf_probe": (ci, [vp, vp, C.c_longlong, vp, vp]),
    "orbx_set_option": (ci, [vp, C.c_char_p, ci]),
    "orbx_set_keypoint_policy": (ci, [vp, ci]),
    "orbx_get_level_tables": (ci, [vp, vp, vp, vp, vp, vp, vp]),
    "orbx_get_level_size": (ci, [vp, ci, vp, vp]),
    "orbx_compute_level_tables": (ci, [ci, cf, ci, vp, vp, vp, vp, vp, vp]),
    "orbx_extract": (ci, [vp, vp, sz, vp, vp, vp, ci, vp, vp]),
    "orbx_frame_construct": (ci, [vp, vp, sz, vp, vp, vp, ci, vp, vp, vp, vp, vp]),
    "orbx_extract_batch": (ci, [vp, vp, sz, ci, vp, vp, vp, ci, vp, vp]),
    "orbx_get_pyramid_level": (ci, [vp, ci, ci, ci, vp, sz]),
    "orbx_get_blurred_level": (ci, [vp, ci, ci, vp, sz]),
    "orbx_get_candidates": (ci, [vp, ci, ci, vp, ci]),
    "orbx_get_level_keypoints": (ci, [vp, ci, ci, vp, vp, ci]),
    "orbx_set_camera": (ci, [vp, vp, vp, vp, vp]),
    "orbx_frame_glue": (ci, [vp, vp, ci, vp, vp, vp]),
    "orbx_features_in_area": (ci, [vp, vp, cf, cf, cf, ci, ci, vp, ci]),
    "orbx_search_for_initialization": (ci, [vp, vp, vp, cf, ci, ci, vp, vp, vp]),
    "orbx_stream_features_in_area": (ci, [vp, ci, vp, vp, vp, ci, vp]),
    "orbx_match_rows": (ci, [vp, vp, vp, ci, ci, ci, vp, vp]),
    "orbx_match_resolve": (ci, [vp, vp, vp, cf, ci, ci, vp, ci, vp, vp, vp]),
    "orbx_hamming_best2": (ci, [vp, vp, ci, vp, ci, vp, vp, vp, ci, vp]),
    "orbx_hamming_best2_noalloc": (ci, [vp, vp, ci, vp, ci, vp, vp, vp, vp, ci]),
    "orbx_descriptor_distance": (ci, [vp, vp]),
    "orbx_descriptor_distance_batch": (ci, [vp, vp, vp, ci, vp]),
    "orbx_check_homography": (ci, [vp, vp, ci, vp, ci, vp, vp, vp, ci, cf, vp, vp, vp, vp]),
    "orbx_check_fundamental": (ci, [vp, vp, ci, vp, ci, vp, vp, ci, cf, vp, vp, vp, vp]),
    "orbx_stream_score_models": (ci, [vp, ci, vp, vp, vp, ci, cf, vp, vp, vp, vp, vp, vp, ci, vp]),
    "orbx_stream_upload": (ci, [vp, vp, ci]),
    "orbx_stream_upload_color": (ci, [vp, vp, ci, ci]),
    "orbx_stream_run": (ci, [vp, ci, cf, ci, ci]),
    "orbx_stream_download": (ci, [vp, ci, vp, vp, vp, ci, vp, vp, vp, vp]),
    "orbx_stream_process": (ci, [vp, vp, ci, cf, ci, ci, ci, vp, vp, vp, ci, vp, vp, vp, vp]),
    "orbx_stream_process_async": (ci, [vp, vp, ci, cf, ci, ci, ci, vp, vp, vp, ci, vp, vp, vp, vp]),
    "orbx_stream_wait": (ci, [vp, ci]),
    "orbx_stream_sync": (ci, [vp]),
    "orbx_timer_start": (ci, [vp]),
    "orbx_timer_stop": (ci, [vp, vp]),
    "orbx_set_stage_timing": (ci, [vp, ci]),
    "orbx_get_stage_times": (ci, [vp, vp]),
    "orbx_last_launch_count": (ci, [vp]),
    "orbx_hostlink_probe": (ci, [vp, vp, sz, vp, sz, ci, vp]),
    "orbx_host_alloc": (vp, [sz]),
    "orbx_host_alloc_ex": (vp, [sz, ci]),
    "orbx_host_free": (None, [vp]),
    "orbx_get_stream": (vp, [vp]),
}

_LIB = None


def lib():
    """The loaded liborbx.so.  Raises if the CUDA extension has not been built."""
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            raise ImportError("%s is missing: build it with `python -m orb_slam_tracking_b200.build` "
                              "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)  # AttributeError if a declared symbol is not exported
            f.restype, f.argtypes = res, args
        _LIB = L
    return _LIB


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def check(rc, ctx=None):
    if rc < 0:
        msg = ""
        if ctx:
            msg = (lib().orbx_last_error(ctx) or b"").decode()
        raise OrbxError(rc, msg)
    return rc
