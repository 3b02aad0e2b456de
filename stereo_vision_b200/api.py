"""cv2-compatible call surface for the dense stereo hot path.

Drop-in for the cv2 names the reference's helpers use or would use:
    StereoSGBM_create(...).compute(left, right)      (north_star; Q made at checkerboard_distance.py:49-51)
    initUndistortRectifyMap / remap                  (checkerboard_distance.py:54-57, 71-72;
                                                      calibrate_v4/distance.py:148-155; calibrate_v4/extrinsic.py:371-378)
    reprojectImageTo3D, filterSpeckles, medianBlur(ksize=3, int16)
    box_median_depth                                 (dense re-definition of ball_drop.py:89-160 + :225-259)

NumPy in -> NumPy out (H2D/D2H inside).  torch CUDA tensor in -> torch CUDA tensor out, zero-copy on the current
stream.  Every function runs hand-written sm_100a kernels through the C ABI (include/svb.h); there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import SgbmParams, SgbmGeometry, check

STEREO_SGBM_MODE_SGBM = 0
STEREO_SGBM_MODE_HH = 1
STEREO_SGBM_MODE_SGBM_3WAY = 2
STEREO_SGBM_MODE_HH4 = 3
INTER_LINEAR = 1
BORDER_CONSTANT = 0
CV_32FC1 = 5
CV_16SC2 = 11

_PARAM_NAMES = ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
                "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")


def _device_of(*xs) -> int:
    for x in xs:
        if isinstance(x, torch.Tensor):
            if not x.is_cuda:
                raise ValueError("torch tensors must live on a CUDA device (no CPU path)")
            return x.device.index
    if not torch.cuda.is_available():
        raise _lib.SvbError("no CUDA device available — stereo_vision_b200 has no CPU fallback")
    return torch.cuda.current_device()


def _stream(dev: int) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _as_dev(x, dev: int, dtype=None) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        t = x
    else:
        a = np.ascontiguousarray(x)
        t = torch.from_numpy(a).to(f"cuda:{dev}", non_blocking=False)
    if dtype is not None and t.dtype != dtype:
        raise ValueError(f"expected dtype {dtype}, got {t.dtype}")
    return t.contiguous()


def _ptr(t) -> C.c_void_p:
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


_ws_cache: dict = {}


def _workspace(dev: int, nbytes: int) -> torch.Tensor:
    t = _ws_cache.get(dev)
    if t is None or t.numel() < nbytes:
        _ws_cache.pop(dev, None)
        t = None
        torch.cuda.synchronize(dev)
        t = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{dev}")
        _ws_cache[dev] = t
    return t


def release_workspaces() -> None:
    _ws_cache.clear()


def _out(t: torch.Tensor, like):
    return t if isinstance(like, torch.Tensor) else t.cpu().numpy()


class StereoSGBM:
    """Mirror of cv2.StereoSGBM: same 11 parameters, getters/setters and compute()."""

    def __init__(self, minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0, preFilterCap=0,
                 uniquenessRatio=0, speckleWindowSize=0, speckleRange=0, mode=STEREO_SGBM_MODE_SGBM):
        vals = (minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff, preFilterCap, uniquenessRatio,
                speckleWindowSize, speckleRange, mode)
        self._p = {n: int(v) for n, v in zip(_PARAM_NAMES, vals)}

    # ---- parameters ----
    def _raw(self) -> SgbmParams:
        return SgbmParams(*[self._p[n] for n in _PARAM_NAMES])

    def effective_params(self) -> dict:
        out = SgbmParams()
        check(_lib.load().svb_sgbm_resolve_params(C.byref(self._raw()), C.byref(out)))
        return {n: getattr(out, n) for n in _PARAM_NAMES}

    def geometry(self, width: int, channels: int = 1) -> dict:
        g = SgbmGeometry()
        check(_lib.load().svb_sgbm_get_geometry(C.byref(self._raw()), int(width), int(channels), C.byref(g)))
        return {n: getattr(g, n) for n, _ in SgbmGeometry._fields_}

    # ---- compute ----
    @staticmethod
    def _check_pair(left, right, batched: bool):
        ls, rs = tuple(left.shape), tuple(right.shape)
        ld = left.dtype
        if ls != rs or ld != right.dtype or ld not in (np.uint8, torch.uint8):
            raise ValueError("left.size() == right.size() && left.type() == right.type() && left.depth() == CV_8U")
        nd = len(ls) - (1 if batched else 0)
        if nd == 2:
            ch = 1
        elif nd == 3 and ls[-1] in (1, 3):
            ch = ls[-1]
        else:
            raise ValueError(f"images must be HxW or HxWx{{1,3}} uint8 (got shape {ls})")
        return ch

    def compute(self, left, right):
        """left/right: HxW or HxWx3 uint8 (NumPy or torch CUDA).  Returns int16 HxW disparity * 16."""
        self._check_pair(left, right, False)
        dev = _device_of(left, right)
        L = _as_dev(left, dev, torch.uint8)
        R = _as_dev(right, dev, torch.uint8)
        out = self._run(L[None], R[None], dev)[0]
        return _out(out, left)

    def compute_batch(self, left, right):
        """Extension: [B,H,W] or [B,H,W,3] uint8 -> [B,H,W] int16.  All pairs are in flight concurrently."""
        self._check_pair(left, right, True)
        dev = _device_of(left, right)
        out = self._run(_as_dev(left, dev, torch.uint8), _as_dev(right, dev, torch.uint8), dev)
        return _out(out, left)

    def _run(self, L: torch.Tensor, R: torch.Tensor, dev: int, out: torch.Tensor | None = None) -> torch.Tensor:
        lib = _lib.load()
        B, H, W = L.shape[0], L.shape[1], L.shape[2]
        ch = 1 if L.dim() == 3 else L.shape[3]
        raw = self._raw()
        nb = C.c_size_t()
        check(lib.svb_sgbm_workspace_bytes(C.byref(raw), H, W, ch, B, C.byref(nb)))
        with torch.cuda.device(dev):
            ws = _workspace(dev, nb.value)
            if out is None:
                out = torch.empty((B, H, W), dtype=torch.int16, device=L.device)
            check(lib.svb_sgbm_compute(_lib.context(dev), C.byref(raw), _ptr(L), _ptr(R), H, W, ch, W * ch, H * W * ch, B,
                                       _ptr(out), W, H * W, _ptr(ws), ws.numel(), _stream(dev)))
        return out

    def debug_stages(self, left, right):
        """Stage taps for parity tests: (C [H,W1,Dp], S_pass1 [H,W1,Dp], disp_raw [H,W]) as NumPy."""
        ch = self._check_pair(left, right, False)
        dev = _device_of(left, right)
        L = _as_dev(left, dev, torch.uint8)
        R = _as_dev(right, dev, torch.uint8)
        H, W = L.shape[0], L.shape[1]
        g = self.geometry(W, ch)
        Cv = torch.empty((H, g["width1"], g["Dp"]), dtype=torch.int16, device=L.device)
        Sv = torch.empty_like(Cv)
        raw_d = torch.empty((H, W), dtype=torch.int16, device=L.device)
        with torch.cuda.device(dev):
            check(_lib.load().svb_sgbm_debug_stages(_lib.context(dev), C.byref(self._raw()), _ptr(L), _ptr(R), H, W, ch,
                                                    W * ch, _ptr(Cv), _ptr(Sv), _ptr(raw_d), _stream(dev)))
        return Cv.cpu().numpy(), Sv.cpu().numpy(), raw_d.cpu().numpy()


def _mk_accessors():
    for n in _PARAM_NAMES:
        cap = n[0].upper() + n[1:]

        def getter(self, _n=n):
            return self._p[_n]

        def setter(self, v, _n=n):
            self._p[_n] = int(v)

        setattr(StereoSGBM, "get" + cap, getter)
        setattr(StereoSGBM, "set" + cap, setter)


_mk_accessors()


def StereoSGBM_create(minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0, preFilterCap=0,
                      uniquenessRatio=0, speckleWindowSize=0, speckleRange=0, mode=STEREO_SGBM_MODE_SGBM) -> StereoSGBM:
    return StereoSGBM(minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff, preFilterCap, uniquenessRatio,
                      speckleWindowSize, speckleRange, mode)


# ---------------------------------------------------------------- rectification
def initUndistortRectifyMap(cameraMatrix, distCoeffs, R, newCameraMatrix, size, m1type=CV_32FC1, device=None):
    """cv2.initUndistortRectifyMap(K, dist, R, P, (W,H), m1type) -> (map1, map2): float32 HxW each for CV_32FC1 (the
    form the reference uses); int16 HxWx2 + uint16 HxW for CV_16SC2 (the form cv2.undistort uses internally).
    Returns NumPy maps unless `device` is given (then torch CUDA tensors that stay resident)."""
    if m1type not in (CV_32FC1, CV_16SC2):
        raise NotImplementedError("only m1type=CV_32FC1 or CV_16SC2 is implemented")
    fixed = m1type == CV_16SC2
    W, H = int(size[0]), int(size[1])
    K = np.ascontiguousarray(np.asarray(cameraMatrix, np.float64).reshape(3, 3))
    dist = np.zeros(0) if distCoeffs is None else np.ascontiguousarray(np.asarray(distCoeffs, np.float64).ravel())
    Rm = None if R is None else np.ascontiguousarray(np.asarray(R, np.float64).reshape(3, 3))
    P33 = None
    if newCameraMatrix is not None:
        P = np.asarray(newCameraMatrix, np.float64)
        if P.shape not in ((3, 3), (3, 4)):
            raise ValueError("newCameraMatrix must be 3x3 or 3x4")
        P33 = np.ascontiguousarray(P[:3, :3])
    keep_dev = device is not None
    dev = _device_of() if device is None else torch.device(device).index or 0
    with torch.cuda.device(dev):
        if fixed:
            mx = torch.empty((H, W, 2), dtype=torch.int16, device=f"cuda:{dev}")
            my = torch.empty((H, W), dtype=torch.uint16, device=f"cuda:{dev}")
        else:
            mx = torch.empty((H, W), dtype=torch.float32, device=f"cuda:{dev}")
            my = torch.empty_like(mx)
        fn = _lib.load().svb_init_undistort_rectify_map_fixed if fixed else _lib.load().svb_init_undistort_rectify_map
        check(fn(
            _lib.context(dev), K.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p) if dist.size else None,
            int(dist.size), Rm.ctypes.data_as(C.c_void_p) if Rm is not None else None,
            P33.ctypes.data_as(C.c_void_p) if P33 is not None else None, W, H, _ptr(mx), _ptr(my), _stream(dev)))
    if keep_dev:
        return mx, my
    return mx.cpu().numpy(), my.cpu().numpy()


def remap(src, map1, map2, interpolation=INTER_LINEAR, dst=None, borderMode=BORDER_CONSTANT, borderValue=0):
    """cv2.remap(src, map1, map2, cv2.INTER_LINEAR) for uint8 HxW / HxWx3 (or batched [B,H,W(,3)] tensors with
    shared maps); float32 HxW maps or the CV_16SC2 pair (int16 HxWx2, uint16 HxW); BORDER_CONSTANT 0.  Bit-exact with
    cv2's fixed-point tables."""
    if interpolation != INTER_LINEAR or borderMode != BORDER_CONSTANT or np.any(np.asarray(borderValue) != 0):
        raise NotImplementedError("only INTER_LINEAR with BORDER_CONSTANT(0) (the reference's call) is implemented")
    if src.dtype not in (np.uint8, torch.uint8):
        raise ValueError("remap: only uint8 images are supported")
    dev = _device_of(src, map1, map2)
    S = _as_dev(src, dev, torch.uint8)
    fixed = map1.dtype in (np.int16, torch.int16)
    if fixed:
        mx = _as_dev(map1, dev, torch.int16)
        my = _as_dev(map2, dev, torch.uint16)
        if mx.dim() != 3 or mx.shape[2] != 2 or tuple(mx.shape[:2]) != tuple(my.shape):
            raise ValueError("fixed-point maps must be int16 HxWx2 and uint16 HxW")
    else:
        mx = _as_dev(map1, dev, torch.float32)
        my = _as_dev(map2, dev, torch.float32)
        if mx.shape != my.shape or mx.dim() != 2:
            raise ValueError("map1/map2 must be float32 HxW arrays of equal shape")
    dH, dW = my.shape
    batched = isinstance(src, torch.Tensor) and (S.dim() == 4 or (S.dim() == 3 and S.shape[-1] not in (1, 3)))
    Sb = S if batched else S[None]
    ch = 1 if Sb.dim() == 3 else Sb.shape[3]
    if ch not in (1, 3):
        raise ValueError("remap: 1- or 3-channel images only")
    B, sH, sW = Sb.shape[0], Sb.shape[1], Sb.shape[2]
    shape = (B, dH, dW) if Sb.dim() == 3 else (B, dH, dW, ch)
    with torch.cuda.device(dev):
        out = torch.empty(shape, dtype=torch.uint8, device=S.device)
        fn = _lib.load().svb_remap_u8_fixed if fixed else _lib.load().svb_remap_u8
        check(fn(_lib.context(dev), _ptr(Sb), sH, sW, ch, sW * ch, sH * sW * ch, _ptr(mx), _ptr(my),
                 _ptr(out), dH, dW, dW * ch, dH * dW * ch, B, _stream(dev)))
    out = out if batched else out[0]
    return _out(out, src)


# ---------------------------------------------------------------- post filters
def medianBlur(src, ksize):
    """cv2.medianBlur(disp_int16, 3) (BORDER_REPLICATE) — the filter compute() applies internally."""
    if ksize != 3 or src.dtype not in (np.int16, torch.int16) or src.ndim != 2:
        raise NotImplementedError("only ksize=3 on int16 HxW maps is implemented")
    dev = _device_of(src)
    s = _as_dev(src, dev, torch.int16)
    H, W = s.shape
    with torch.cuda.device(dev):
        out = torch.empty_like(s)
        check(_lib.load().svb_median3_s16(_lib.context(dev), _ptr(s), _ptr(out), H, W, W, 1, H * W, _stream(dev)))
    return _out(out, src)


def filterSpeckles(img, newVal, maxSpeckleSize, maxDiff, buf=None):
    """cv2.filterSpeckles(img, newVal, maxSpeckleSize, maxDiff): in place on int16 HxW; returns (img, None)."""
    if img.dtype not in (np.int16, torch.int16) or img.ndim != 2:
        raise ValueError("filterSpeckles: int16 HxW image required")
    dev = _device_of(img)
    t = _as_dev(img, dev, torch.int16)
    if not isinstance(img, torch.Tensor):
        t = t.clone()
    H, W = t.shape
    with torch.cuda.device(dev):
        labels = torch.empty(2 * H * W, dtype=torch.int32, device=t.device)
        check(_lib.load().svb_filter_speckles_s16(_lib.context(dev), _ptr(t), H, W, W, 1, H * W, int(newVal),
                                                  int(maxSpeckleSize), int(maxDiff), _ptr(labels), labels.numel(),
                                                  _stream(dev)))
    if isinstance(img, torch.Tensor):
        if t.data_ptr() != img.data_ptr():
            img.copy_(t)
        return img, None
    img[...] = t.cpu().numpy()
    return img, None


# ---------------------------------------------------------------- 3-D
_DT = {torch.uint8: 0, torch.int16: 1, torch.int32: 2, torch.float32: 3}


def reprojectImageTo3D(disparity, Q, _3dImage=None, handleMissingValues=False, ddepth=-1):
    """cv2.reprojectImageTo3D(disp, Q, handleMissingValues) -> float32 HxWx3 (raw disparity values, no /16)."""
    if ddepth not in (-1, CV_32FC1):
        raise NotImplementedError("only float32 output (ddepth=-1) is implemented")
    dev = _device_of(disparity)
    d = _as_dev(disparity, dev)
    if d.dtype not in _DT or d.dim() not in (2, 3):
        raise ValueError("disparity must be CV_8UC1, CV_16SC1, CV_32SC1 or CV_32FC1 (HxW)")
    Qh = np.ascontiguousarray(np.asarray(Q, np.float64))
    if Qh.shape != (4, 4):
        raise ValueError("Q must be 4x4")
    db = d if d.dim() == 3 else d[None]
    B, H, W = db.shape
    with torch.cuda.device(dev):
        out = torch.empty((B, H, W, 3), dtype=torch.float32, device=d.device)
        mins = torch.empty(B, dtype=torch.float64, device=d.device) if handleMissingValues else None
        check(_lib.load().svb_reproject_to_3d(_lib.context(dev), _ptr(db), _DT[d.dtype], H, W,
                                              Qh.ctypes.data_as(C.c_void_p), int(bool(handleMissingValues)), _ptr(out),
                                              B, _ptr(mins), _stream(dev)))
    out = out if d.dim() == 3 else out[0]
    return _out(out, disparity)


def box_median_depth(disp16, Q, boxes, minDisparity=0, box_image=None):
    """Fused reprojection + per-box median depth.  disp16: int16 HxW (or [B,H,W]) as compute() returns;
    boxes: N x 4 float32 xyxy (ultralytics boxes.data[:, :4]) in left-image pixels; box_image: frame index per box
    for batched maps.  Returns float32 N: median Z over valid pixels (NaN when none).  SURVEY.md §8a row a10."""
    dev = _device_of(disp16, boxes)
    d = _as_dev(disp16, dev, torch.int16)
    bx = _as_dev(np.asarray(boxes, np.float32).reshape(-1, 4) if not isinstance(boxes, torch.Tensor) else boxes, dev,
                 torch.float32).reshape(-1, 4)
    Qh = np.ascontiguousarray(np.asarray(Q, np.float64))
    if Qh.shape != (4, 4):
        raise ValueError("Q must be 4x4")
    db = d if d.dim() == 3 else d[None]
    _, H, W = db.shape
    bi = None
    if box_image is not None:
        bi = _as_dev(np.asarray(box_image, np.int32) if not isinstance(box_image, torch.Tensor) else box_image, dev,
                     torch.int32)
    n = bx.shape[0]
    with torch.cuda.device(dev):
        out = torch.empty(n, dtype=torch.float32, device=d.device)
        check(_lib.load().svb_box_median_depth(_lib.context(dev), _ptr(db), H, W, H * W, Qh.ctypes.data_as(C.c_void_p),
                                               (int(minDisparity) - 1) * 16, _ptr(bx), _ptr(bi), n, _ptr(out),
                                               _stream(dev)))
    return out if isinstance(disp16, torch.Tensor) else out.cpu().numpy()


# ---------------------------------------------------------------- measurement hooks
STAGE_NAMES = ("planes", "cost", "aggregate_pass1", "aggregate_pass2_select", "lrcheck_median", "speckle")


def set_latency_form(enable: bool, device=None) -> None:
    """Small jobs (a single pair, e.g. the drop-in compute(L, R) call) run their two sweeps concurrently by default;
    False forces the back-to-back form that large batches use (identical results; see include/svb.h)."""
    dev = torch.cuda.current_device() if device is None else int(device)
    check(_lib.load().svb_set_latency_form(_lib.context(dev), int(bool(enable))))


def set_stage_timing(enable: bool, device=None) -> None:
    """Record CUDA events between the stages of StereoSGBM.compute on this device's context."""
    dev = _device_of() if device is None else int(device)
    check(_lib.load().svb_set_timing(_lib.context(dev), int(bool(enable))))


def last_stage_ms(device=None) -> dict:
    """Per-stage device milliseconds of the last compute() (blocks until it has finished)."""
    dev = _device_of() if device is None else int(device)
    ms = (C.c_float * 6)()
    check(_lib.load().svb_sgbm_last_stage_ms(_lib.context(dev), ms, 6))
    return dict(zip(STAGE_NAMES, [float(v) for v in ms]))


def packed_alu_peak(device=None) -> dict:
    """Measured device peak of packed s16x2 ALU instructions (svb_measure_packed_alu_peak): the roofline denominator
    of the aggregation stage.  Returns packed instructions/s (thread level) and 16-bit lane operations/s (x2)."""
    dev = torch.cuda.current_device() if device is None else int(device)
    rate, mhz = C.c_double(), C.c_double()
    check(_lib.load().svb_measure_packed_alu_peak(_lib.context(dev), C.byref(rate), C.byref(mhz), _stream(dev)))
    return {"packed_instr_per_s": rate.value, "ops16_per_s": 2.0 * rate.value, "sm_clock_max_mhz": mhz.value}


# ---------------------------------------------------------------- sparse 3-D path + ingest
COLOR_BGR2GRAY = 6


def _host_f64(a, shape=None):
    if a is None:
        return None
    a = np.ascontiguousarray(np.asarray(a, np.float64))
    return a.reshape(shape) if shape is not None else a


def _hp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def undistortPoints(src, cameraMatrix, distCoeffs, R=None, P=None):
    """cv2.undistortPoints(src, K, dist, R=None, P=None): float32/float64 points shaped (N,1,2), (1,N,2) or (N,2);
    same dtype out, always shaped (N,1,2) as cv2 returns it (ball_drop.py:239-240, 3dpose.py:538-549, checkerboard_distance.py:83-86)."""
    is_t = isinstance(src, torch.Tensor)
    dt = src.dtype
    if dt not in (np.float32, np.float64, torch.float32, torch.float64):
        raise ValueError("undistortPoints: points must be float32 or float64")
    if src.shape[-1] != 2 or src.ndim not in (2, 3):
        raise ValueError("undistortPoints: points must be shaped (N,1,2), (1,N,2) or (N,2)")
    dev = _device_of(src)
    S = _as_dev(src, dev)
    n = S.numel() // 2
    K = _host_f64(cameraMatrix, (3, 3))
    dist = np.zeros(0) if distCoeffs is None else _host_f64(distCoeffs).ravel()
    Rm = _host_f64(R, (3, 3)) if R is not None else None
    P33 = None
    if P is not None:
        Pm = np.asarray(P, np.float64)
        if Pm.shape not in ((3, 3), (3, 4)):
            raise ValueError("P must be 3x3 or 3x4")
        P33 = np.ascontiguousarray(Pm[:3, :3])
    with torch.cuda.device(dev):
        out = torch.empty_like(S)
        code = 3 if S.dtype == torch.float32 else 4
        check(_lib.load().svb_undistort_points(_lib.context(dev), _ptr(S), _ptr(out), code, n, _hp(K), _hp(dist),
                                               int(dist.size), _hp(Rm), _hp(P33), _stream(dev)))
    out = out.reshape(n, 1, 2)
    return out if is_t else out.cpu().numpy()


def triangulatePoints(projMatr1, projMatr2, projPoints1, projPoints2):
    """cv2.triangulatePoints(P1, P2, pts1, pts2) -> 4xN homogeneous points (float32 for float32 points, else float64).
    Points: 2xN (cv2's documented layout), (N,1,2) / (1,N,2) (what undistortPoints returns) or a single (2,) point
    (ball_drop.py:254).  The vector is unit-norm with arbitrary sign, as with cv2: divide by row 3."""
    is_t = isinstance(projPoints1, torch.Tensor)

    def norm(p):
        if not isinstance(p, torch.Tensor):
            p = np.asarray(p)
            if p.dtype not in (np.float32, np.float64):
                p = p.astype(np.float64)
        if p.ndim == 1 and p.shape[0] == 2:
            return p.reshape(1, 2), 1
        if p.ndim == 3 and p.shape[-1] == 2:
            return p.reshape(-1, 2), 1
        if p.ndim == 2 and p.shape[0] == 2:
            return p, 0
        if p.ndim == 2 and p.shape[1] == 2:
            return p, 1
        raise ValueError("triangulatePoints: points must be 2xN, Nx2, (N,1,2), (1,N,2) or a single (2,) point")
    a, la = norm(projPoints1)
    b, lb = norm(projPoints2)
    dev = _device_of(a, b)
    A, Bm = _as_dev(a, dev), _as_dev(b, dev)
    if lb != la:
        Bm = Bm.t().contiguous()
    if A.dtype != Bm.dtype:
        A, Bm = A.double(), Bm.double()
    n = A.shape[1] if la == 0 else A.shape[0]
    if (Bm.shape[1] if la == 0 else Bm.shape[0]) != n:
        raise ValueError("triangulatePoints: both views need the same number of points")
    P1 = _host_f64(projMatr1, (3, 4))
    P2 = _host_f64(projMatr2, (3, 4))
    with torch.cuda.device(dev):
        out = torch.empty((4, n), dtype=A.dtype, device=A.device)
        code = 3 if A.dtype == torch.float32 else 4
        check(_lib.load().svb_triangulate_points(_lib.context(dev), _hp(P1), _hp(P2), _ptr(A), _ptr(Bm), code, n, la,
                                                 _ptr(out), _stream(dev)))
    return out if is_t else out.cpu().numpy()


def cvtColor(src, code):
    """cv2.cvtColor(img, cv2.COLOR_BGR2GRAY) for uint8 HxWx3 (or batched [B,H,W,3] tensors): cv2's 15-bit fixed point."""
    if code != COLOR_BGR2GRAY:
        raise NotImplementedError("only COLOR_BGR2GRAY is implemented")
    if src.dtype not in (np.uint8, torch.uint8) or src.shape[-1] != 3 or src.ndim not in (3, 4):
        raise ValueError("cvtColor: uint8 HxWx3 image expected")
    dev = _device_of(src)
    S = _as_dev(src, dev, torch.uint8)
    Sb = S if S.dim() == 4 else S[None]
    B, H, W = Sb.shape[0], Sb.shape[1], Sb.shape[2]
    with torch.cuda.device(dev):
        out = torch.empty((B, H, W), dtype=torch.uint8, device=S.device)
        check(_lib.load().svb_bgr2gray_u8(_lib.context(dev), _ptr(Sb), _ptr(out), H, W, 3 * W, 3 * H * W, W, H * W, B,
                                          _stream(dev)))
    out = out if S.dim() == 4 else out[0]
    return _out(out, src)


def undistort(src, cameraMatrix, distCoeffs, dst=None, newCameraMatrix=None):
    """cv2.undistort(img, K, dist[, None, newK]) = initUndistortRectifyMap(K, dist, I, newK or K, CV_16SC2) +
    remap(INTER_LINEAR), bit-exact (calibrate_v4/distance.py:125-126).  uint8 HxW or HxWx3."""
    if src.ndim not in (2, 3):
        raise ValueError("undistort: HxW or HxWx3 image expected")
    H, W = src.shape[:2]
    dev = _device_of(src)
    mx, my = initUndistortRectifyMap(cameraMatrix, distCoeffs, None, cameraMatrix if newCameraMatrix is None else newCameraMatrix,
                                     (int(W), int(H)), CV_16SC2, device=f"cuda:{dev}")
    return remap(src, mx, my, INTER_LINEAR)


def triangulate_stereo(left_pts, right_pts, K1, d1, K2, d2, R, T):
    """Batched form of the reference's calculate_3d_position (ball_drop.py:225-259): undistortPoints(P=K) on both views,
    P1 = K1[I|0], P2 = K2[R|T], triangulatePoints, divide by w.  left/right points: (N,2) pixels.  Returns (N,3)."""
    lp = np.asarray(left_pts, np.float32).reshape(-1, 1, 2) if not isinstance(left_pts, torch.Tensor) else left_pts
    rp = np.asarray(right_pts, np.float32).reshape(-1, 1, 2) if not isinstance(right_pts, torch.Tensor) else right_pts
    lu = undistortPoints(lp, K1, d1, P=K1)
    ru = undistortPoints(rp, K2, d2, P=K2)
    K1 = np.asarray(K1, np.float64); K2 = np.asarray(K2, np.float64)
    P1 = K1 @ np.hstack((np.eye(3), np.zeros((3, 1))))
    P2 = K2 @ np.hstack((np.asarray(R, np.float64).reshape(3, 3), np.asarray(T, np.float64).reshape(3, 1)))
    X = triangulatePoints(P1, P2, lu, ru)
    return (X[:3] / X[3]).T
