"""ctypes binding of libsvb.so (include/svb.h).  No CPU fallback: if the library or a CUDA device is missing,
every compute entry raises."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libsvb.so")


class SvbError(RuntimeError):
    pass


class SgbmParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


class SgbmGeometry(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("minX1", "maxX1", "width1", "D", "Dp", "invalid16", "max_cost", "safe_packed")]


# every symbol include/svb.h declares (tests check that the library exports all of them)
EXPORTS = [
    "svb_last_error", "svb_version", "svb_create", "svb_destroy", "svb_sgbm_resolve_params", "svb_sgbm_get_geometry",
    "svb_sgbm_workspace_bytes", "svb_sgbm_compute", "svb_sgbm_debug_stages", "svb_median3_s16",
    "svb_filter_speckles_s16", "svb_init_undistort_rectify_map", "svb_remap_u8", "svb_reproject_to_3d",
    "svb_box_median_depth", "svb_set_timing", "svb_sgbm_last_stage_ms", "svb_measure_packed_alu_peak",
    "svb_undistort_points", "svb_triangulate_points", "svb_bgr2gray_u8", "svb_init_undistort_rectify_map_fixed",
    "svb_remap_u8_fixed", "svb_filter_speckles_s16_noalloc", "svb_set_latency_form",
]

_lib = None


def load():
    """Load libsvb.so (built by stereo_vision_b200.build).  Raises if it is missing — there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise SvbError(f"{SO_PATH} not found: build it with `python -m stereo_vision_b200.build` "
                       "(the engine has no CPU fallback)")
    lib = C.CDLL(SO_PATH)
    vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int
    lib.svb_last_error.restype = C.c_char_p
    lib.svb_version.restype = C.c_char_p
    lib.svb_create.argtypes = [i32, C.POINTER(vp)]
    lib.svb_destroy.argtypes = [vp]
    lib.svb_sgbm_resolve_params.argtypes = [C.POINTER(SgbmParams), C.POINTER(SgbmParams)]
    lib.svb_sgbm_get_geometry.argtypes = [C.POINTER(SgbmParams), i32, i32, C.POINTER(SgbmGeometry)]
    lib.svb_sgbm_workspace_bytes.argtypes = [C.POINTER(SgbmParams), i32, i32, i32, i32, C.POINTER(sz)]
    lib.svb_sgbm_compute.argtypes = [vp, C.POINTER(SgbmParams), vp, vp, i32, i32, i32, sz, sz, i32, vp, sz, sz, vp, sz, vp]
    lib.svb_sgbm_debug_stages.argtypes = [vp, C.POINTER(SgbmParams), vp, vp, i32, i32, i32, sz, vp, vp, vp, vp]
    lib.svb_median3_s16.argtypes = [vp, vp, vp, i32, i32, sz, i32, sz, vp]
    lib.svb_filter_speckles_s16.argtypes = [vp, vp, i32, i32, sz, i32, sz, i32, i32, i32, vp, sz, vp]
    lib.svb_filter_speckles_s16_noalloc.argtypes = [vp, vp, i32, i32, sz, i32, sz, i32, i32, i32, vp]
    lib.svb_init_undistort_rectify_map.argtypes = [vp, vp, vp, i32, vp, vp, i32, i32, vp, vp, vp]
    lib.svb_remap_u8.argtypes = [vp, vp, i32, i32, i32, sz, sz, vp, vp, vp, i32, i32, sz, sz, i32, vp]
    lib.svb_init_undistort_rectify_map_fixed.argtypes = lib.svb_init_undistort_rectify_map.argtypes
    lib.svb_remap_u8_fixed.argtypes = lib.svb_remap_u8.argtypes
    lib.svb_reproject_to_3d.argtypes = [vp, vp, i32, i32, i32, vp, i32, vp, i32, vp, vp]
    lib.svb_box_median_depth.argtypes = [vp, vp, i32, i32, sz, vp, i32, vp, vp, i32, vp, vp]
    lib.svb_set_timing.argtypes = [vp, i32]
    lib.svb_set_latency_form.argtypes = [vp, i32]
    lib.svb_sgbm_last_stage_ms.argtypes = [vp, C.POINTER(C.c_float), i32]
    lib.svb_undistort_points.argtypes = [vp, vp, vp, i32, i32, vp, vp, i32, vp, vp, vp]
    lib.svb_triangulate_points.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, vp, vp]
    lib.svb_bgr2gray_u8.argtypes = [vp, vp, vp, i32, i32, sz, sz, sz, sz, i32, vp]
    lib.svb_measure_packed_alu_peak.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), vp]
    for n in EXPORTS:
        f = getattr(lib, n)
        if n not in ("svb_last_error", "svb_version"):
            f.restype = i32
    _lib = lib
    return lib


def check(rc: int):
    """Map C-ABI status codes to Python exceptions mirroring cv2.error cases."""
    if rc == 0:
        return
    msg = load().svb_last_error().decode("utf-8", "replace")
    if rc == -1:
        raise ValueError(msg)
    if rc == -3:
        raise NotImplementedError(msg)
    raise SvbError(f"libsvb error {rc}: {msg}")


_ctx = {}


def context(device: int) -> C.c_void_p:
    """One context per (process, device)."""
    if device not in _ctx:
        h = C.c_void_p()
        check(load().svb_create(int(device), C.byref(h)))
        _ctx[device] = h
    return _ctx[device]
