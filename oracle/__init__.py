"""CPU ORACLE — test infrastructure only (see oracle/sv_oracle.c header).

ctypes loader for the plain-C restatement of the cv2 4.13.0 dense-stereo path.  May be imported
only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (stereo_vision_b200) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsv_oracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "sv_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libsv_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


class Params(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.svo_sgbm_compute.restype = C.c_int
        _lib.svo_sgbm_compute_raw.restype = C.c_int
        _lib.svo_sgbm_cost_volume.restype = C.c_int
        _lib.svo_filter_speckles_s16.restype = C.c_int
        _lib.svo_init_undistort_rectify_map.restype = C.c_int
        _lib.svo_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def make_params(minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0, preFilterCap=0,
                uniquenessRatio=0, speckleWindowSize=0, speckleRange=0, mode=0) -> Params:
    return Params(minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff, preFilterCap, uniquenessRatio,
                  speckleWindowSize, speckleRange, mode)


def resolve(**kw) -> dict:
    pin, pout = make_params(**kw), Params()
    lib().svo_resolve_params(C.byref(pin), C.byref(pout))
    return {n: getattr(pout, n) for n, _ in Params._fields_}


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if a.ndim == 2:
        a = a[:, :, None]
    return a


def geometry(W, **kw):
    r = resolve(**kw)
    minD, D = r["minDisparity"], r["numDisparities"]
    maxD = minD + D
    minX1, maxX1 = max(maxD, 0), W + min(minD, 0)
    return minX1, maxX1, maxX1 - minX1, D


def sgbm_compute(left, right, **kw) -> np.ndarray:
    L, R = _img(left), _img(right)
    assert L.shape == R.shape
    H, W, ch = L.shape
    out = np.empty((H, W), np.int16)
    p = make_params(**kw)
    rc = lib().svo_sgbm_compute(_p(L), _p(R), H, W, ch, C.c_size_t(W * ch), C.byref(p), _p(out))
    if rc:
        raise ValueError(f"oracle sgbm_compute failed rc={rc}")
    return out


def sgbm_stages(left, right, **kw):
    """Returns (disp_raw, Cvol, S): pre-median disparity, cost volume and aggregated volume."""
    L, R = _img(left), _img(right)
    H, W, ch = L.shape
    _, _, W1, D = geometry(W, **kw)
    Cv = np.empty((H, W1, D), np.int16)
    S = np.empty((H, W1, D), np.int16)
    raw = np.empty((H, W), np.int16)
    p = make_params(**kw)
    rc = lib().svo_sgbm_compute_raw(_p(L), _p(R), H, W, ch, C.c_size_t(W * ch), C.byref(p), _p(raw), _p(Cv), _p(S))
    if rc:
        raise ValueError(f"oracle sgbm_stages failed rc={rc}")
    return raw, Cv, S


def aggregate(Cvol, P1, P2, ndirs) -> np.ndarray:
    """Saturated sum of the first `ndirs` path directions (order: left, up-left, up, up-right, right, ...)."""
    Cvol = np.ascontiguousarray(Cvol, np.int16)
    H, W1, D = Cvol.shape
    S = np.empty_like(Cvol)
    lib().svo_sgbm_aggregate_dirs(_p(Cvol), H, W1, D, int(P1), int(P2), int(ndirs), _p(S))
    return S


def median3(a) -> np.ndarray:
    a = np.ascontiguousarray(a, np.int16)
    out = np.empty_like(a)
    lib().svo_median3_s16(_p(a), a.shape[0], a.shape[1], _p(out))
    return out


def filter_speckles(a, newVal, maxSpeckleSize, maxDiff) -> np.ndarray:
    a = np.array(a, dtype=np.int16, order="C", copy=True)
    lib().svo_filter_speckles_s16(_p(a), a.shape[0], a.shape[1], int(newVal), int(maxSpeckleSize), int(maxDiff))
    return a


def remap(src, mapx, mapy) -> np.ndarray:
    s = _img(src)
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    sH, sW, ch = s.shape
    dH, dW = mapx.shape
    dst = np.empty((dH, dW, ch), np.uint8)
    lib().svo_remap_u8(_p(s), sH, sW, ch, C.c_size_t(sW * ch), _p(mapx), _p(mapy), dH, dW, _p(dst),
                       C.c_size_t(dW * ch))
    return dst if np.ndim(src) == 3 else dst[:, :, 0]


def init_undistort_rectify_map(K, dist, R, P, size):
    W, H = size
    K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
    dist = np.zeros(0) if dist is None else np.ascontiguousarray(dist, np.float64).ravel()
    Rm = None if R is None else np.ascontiguousarray(R, np.float64).reshape(3, 3)
    P33 = None if P is None else np.ascontiguousarray(np.asarray(P, np.float64)[:3, :3])
    mx = np.empty((H, W), np.float32)
    my = np.empty((H, W), np.float32)
    rc = lib().svo_init_undistort_rectify_map(_p(K), _p(dist) if dist.size else None, int(dist.size),
                                              _p(Rm) if Rm is not None else None,
                                              _p(P33) if P33 is not None else None, W, H, _p(mx), _p(my))
    if rc:
        raise ValueError("singular P*R")
    return mx, my


_DT = {np.dtype(np.uint8): 0, np.dtype(np.int16): 1, np.dtype(np.int32): 2, np.dtype(np.float32): 3}


def reproject_to_3d(disp, Q, handleMissingValues=False) -> np.ndarray:
    disp = np.ascontiguousarray(disp)
    Q = np.ascontiguousarray(Q, np.float64).reshape(4, 4)
    H, W = disp.shape
    out = np.empty((H, W, 3), np.float32)
    lib().svo_reproject_to_3d(_p(disp), _DT[disp.dtype], H, W, _p(Q), int(bool(handleMissingValues)), _p(out))
    return out


def box_median_depth(disp16, Q, boxes, invalid16) -> np.ndarray:
    disp16 = np.ascontiguousarray(disp16, np.int16)
    Q = np.ascontiguousarray(Q, np.float64).reshape(4, 4)
    boxes = np.ascontiguousarray(boxes, np.float32).reshape(-1, 4)
    out = np.empty(len(boxes), np.float32)
    lib().svo_box_median_depth(_p(disp16), disp16.shape[0], disp16.shape[1], _p(Q), int(invalid16), _p(boxes),
                               len(boxes), _p(out))
    return out


def num_threads() -> int:
    return lib().svo_num_threads()


def set_num_threads(n: int) -> None:
    lib().svo_set_num_threads(int(n))


# ---- sparse 3-D path: undistortPoints / triangulatePoints / BGR2GRAY ----
def undistort_points(src, K, dist, R=None, P=None) -> np.ndarray:
    """f64 restatement of cv2.undistortPoints; returns float64 (N,2)."""
    p = np.ascontiguousarray(np.asarray(src, dtype=np.float64).reshape(-1, 2))
    K = np.ascontiguousarray(K, dtype=np.float64)
    d = np.ascontiguousarray(np.asarray(dist, dtype=np.float64).ravel()) if dist is not None else np.zeros(0)
    Rm = np.ascontiguousarray(R, dtype=np.float64) if R is not None else None
    Pm = np.ascontiguousarray(np.asarray(P, dtype=np.float64)[:3, :3]) if P is not None else None
    out = np.empty_like(p)
    lib().svo_undistort_points(_p(p), C.c_int(len(p)), _p(K), _p(d), C.c_int(d.size), _p(Rm) if Rm is not None else None,
                               _p(Pm) if Pm is not None else None, _p(out))
    return out


def triangulate_points(P1, P2, a, b) -> np.ndarray:
    """Homogeneous 4xN (float64) DLT triangulation; a, b are (N,2)."""
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1, 2))
    b = np.ascontiguousarray(np.asarray(b, dtype=np.float64).reshape(-1, 2))
    P1 = np.ascontiguousarray(P1, dtype=np.float64); P2 = np.ascontiguousarray(P2, dtype=np.float64)
    out = np.empty((len(a), 4), dtype=np.float64)
    lib().svo_triangulate_points(_p(P1), _p(P2), _p(a), _p(b), C.c_int(len(a)), _p(out))
    return out.T.copy()


def bgr2gray(img) -> np.ndarray:
    img = np.ascontiguousarray(img, dtype=np.uint8)
    out = np.empty(img.shape[:2], dtype=np.uint8)
    lib().svo_bgr2gray(_p(img), C.c_int(out.size), _p(out))
    return out

