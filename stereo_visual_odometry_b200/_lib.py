"""ctypes binding of the C ABI in include/svo.h (libsvo_b200.so, built in-tree).

There is no Python / CPU fallback: if the shared library is missing, or no CUDA
device is present, every compute entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVO_LIB_PATH") or os.path.join(_HERE, "libsvo_b200.so")   # (override: A/B builds)
CSRC = os.path.join(_HERE, "csrc")


class SvoError(RuntimeError):
    pass


class Params(C.Structure):
    """svo_params (include/svo.h); defaults = include/vo.h:53-127 of the reference."""
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("n_seq", C.c_int),
        ("buckets_along_height", C.c_int), ("buckets_along_width", C.c_int),
        ("bucket_start_row", C.c_int), ("features_per_bucket", C.c_int),
        ("features_threshold", C.c_int), ("pre_matching_threshold", C.c_int),
        ("age_threshold", C.c_int), ("fast_threshold", C.c_int),
        ("win_size", C.c_int), ("max_level", C.c_int), ("lk_max_iters", C.c_int),
        ("lk_epsilon", C.c_float), ("lk_min_eig", C.c_float), ("circular_threshold", C.c_float),
        ("ransac_reproj_error", C.c_float), ("ransac_iterations", C.c_int),
        ("ransac_hypotheses", C.c_int), ("ransac_confidence", C.c_float),
        ("max_translation_norm", C.c_double), ("max_rotation_norm", C.c_double),
        ("ransac_seed", C.c_uint64), ("device", C.c_int), ("channels", C.c_int), ("lookahead", C.c_int),
    ]


class FrameResult(C.Structure):
    """svo_frame_result (include/svo.h)."""
    _fields_ = [
        ("success", C.c_int), ("reason", C.c_int), ("n_bucketed", C.c_int), ("n_matched", C.c_int),
        ("n_inliers", C.c_int), ("retry", C.c_int), ("pnp_ok", C.c_int), ("lk_iters", C.c_int),
        ("R", C.c_double * 9), ("t", C.c_double * 3), ("transform", C.c_double * 16),
    ]


REASONS = {0: "", 1: "first", 2: "few_features", 3: "pnp_failed", 4: "too_big"}

# every symbol include/svo.h declares
SYMBOLS = [
    "svo_version", "svo_last_error", "svo_device_count", "svo_default_params", "svo_create",
    "svo_destroy", "svo_set_projection", "svo_process_frame", "svo_process_frame_batch",
    "svo_process_frame_batch_dev", "svo_submit_frame", "svo_submit_frame_batch", "svo_submit_frames", "svo_collect", "svo_pending", "svo_stream_join", "svo_pipeline_info",
    "svo_synchronize", "svo_get_result", "svo_get_tap", "svo_get_frame_images", "svo_stream",
    "svo_launch_count", "svo_enable_timers", "svo_get_stage_time", "svo_reset_timers",
    "svo_fast_detect", "svo_bucket_filter", "svo_build_pyramid", "svo_lk_track",
    "svo_circular_match", "svo_matching_features", "svo_find_close_points", "svo_compact",
    "svo_triangulate", "svo_camera_to_world", "svo_get_inverse_transform", "svo_fast_detect_noalloc",
]

_lib = None


def build(force: bool = False) -> str:
    """Compile libsvo_b200.so for sm_100a with nvcc (csrc/Makefile)."""
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    r = subprocess.run(["make", "-C", CSRC], capture_output=True, text=True)
    if r.returncode != 0 or not os.path.exists(LIB_PATH):
        raise SvoError("building libsvo_b200.so failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SvoError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, ip, fp, dp, u8p = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_float), C.POINTER(C.c_double), C.c_void_p
    L.svo_version.restype = C.c_char_p
    L.svo_last_error.restype = C.c_char_p
    L.svo_default_params.argtypes = [C.POINTER(Params), C.c_int, C.c_int]
    L.svo_default_params.restype = None
    L.svo_create.argtypes = [C.POINTER(Params)]
    L.svo_create.restype = vp
    L.svo_destroy.argtypes = [vp]
    L.svo_destroy.restype = None
    L.svo_set_projection.argtypes = [vp, fp, fp]
    L.svo_process_frame.argtypes = [vp, u8p, u8p, C.c_int, dp]
    L.svo_process_frame_batch.argtypes = [vp, u8p, u8p, C.c_int, C.c_longlong, C.POINTER(FrameResult)]
    L.svo_process_frame_batch_dev.argtypes = [vp, vp, vp, C.c_int, C.c_longlong, vp]
    L.svo_submit_frame.argtypes = [vp, u8p, u8p, C.c_int]
    L.svo_submit_frame_batch.argtypes = [vp, u8p, u8p, C.c_int, C.c_longlong]
    L.svo_submit_frames.argtypes = [vp, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int]
    L.svo_collect.argtypes = [vp, C.POINTER(FrameResult)]
    L.svo_pending.argtypes = [vp]
    L.svo_stream_join.argtypes = [vp]
    L.svo_pipeline_info.argtypes = [vp, ip]
    L.svo_synchronize.argtypes = [vp]
    L.svo_get_result.argtypes = [vp, C.c_int, C.POINTER(FrameResult)]
    L.svo_get_tap.argtypes = [vp, C.c_int, C.c_int, C.c_int, fp, ip, u8p]
    L.svo_get_frame_images.argtypes = [vp, C.c_int, u8p, u8p, C.c_int]
    L.svo_stream.argtypes = [vp]
    L.svo_stream.restype = vp
    L.svo_launch_count.argtypes = [vp]
    L.svo_launch_count.restype = C.c_longlong
    L.svo_enable_timers.argtypes = [vp, C.c_int]
    L.svo_get_stage_time.argtypes = [vp, C.c_int, dp, C.POINTER(C.c_longlong)]
    L.svo_reset_timers.argtypes = [vp]
    L.svo_fast_detect.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, fp, fp, C.c_int]
    L.svo_fast_detect_noalloc.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, vp, vp]
    L.svo_bucket_filter.argtypes = [fp, ip, ip] + [C.c_int] * 9
    L.svo_build_pyramid.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p, C.c_longlong]
    L.svo_lk_track.argtypes = [C.POINTER(Params), u8p, u8p, C.c_int, fp, C.c_int, fp, u8p]
    L.svo_circular_match.argtypes = [C.POINTER(Params), u8p, u8p, u8p, u8p, C.c_int, fp, C.c_int, fp, fp, fp, u8p]
    L.svo_matching_features.argtypes = [C.POINTER(Params), u8p, u8p, u8p, u8p, C.c_int, fp, ip, ip, ip, C.c_int,
                                        fp, fp, fp, fp]
    L.svo_find_close_points.argtypes = [fp, fp, C.c_int, C.c_float, u8p]
    L.svo_compact.argtypes = [vp, C.c_int, C.c_int, u8p]
    L.svo_triangulate.argtypes = [fp, fp, fp, fp, C.c_int, fp]
    L.svo_camera_to_world.argtypes = [C.POINTER(Params), fp, fp, fp, C.c_int, dp, dp, ip, ip]
    L.svo_get_inverse_transform.argtypes = [dp, dp, dp]
    _lib = L
    return L


def check(rc: int) -> int:
    if rc < 0:
        raise SvoError(f"svo error {rc}: {lib().svo_last_error().decode()}")
    return rc


def ptr(a: np.ndarray, t=None):
    """Pointer to a C-contiguous numpy array (void* for byte buffers)."""
    if t is None:
        return a.ctypes.data_as(C.c_void_p)
    return a.ctypes.data_as(C.POINTER(t))


def default_params(width: int, height: int, **kw) -> Params:
    p = Params()
    lib().svo_default_params(C.byref(p), width, height)
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p
