"""Seeded synthetic stereo inputs.  NumPy only — no cv2, no files."""
from __future__ import annotations

import numpy as np


def _gauss_blur_u8(img: np.ndarray, sigma: float = 1.0) -> np.ndarray:
    """Integer binomial blur ([1,4,6,4,1]/16, separable, sigma~1): exact on every host, so seeded inputs
    and the golden vectors made from them are reproducible bit for bit."""
    k = (1, 4, 6, 4, 1)
    a = np.pad(img.astype(np.int32), ((2, 2), (2, 2)), mode="reflect")
    a = sum(k[i] * a[:, i:i + img.shape[1]] for i in range(5))
    a = sum(k[i] * a[i:i + img.shape[0], :] for i in range(5))
    return ((a + 128) >> 8).astype(np.uint8)


def disparity_field(H: int, W: int, D: int) -> np.ndarray:
    y = np.arange(H, dtype=np.float64)[:, None]
    d = np.floor(0.15 * D + 0.6 * D * y / max(H - 1, 1)).astype(np.int64) * np.ones((1, W), np.int64)
    d[H // 3:2 * H // 3, W // 3:2 * W // 3] = int(np.floor(0.85 * D))
    return d


def stereo_pair(H: int, W: int, D: int, seed: int = 1000, kind: str = "smooth", channels: int = 1):
    """Rectified pair with a piecewise disparity field.  kind: smooth | noise | binary."""
    rng = np.random.default_rng(seed)
    outs = []
    for _ in range(channels):
        if kind == "smooth":
            base = _gauss_blur_u8(rng.integers(0, 256, (H, W + D + 8), dtype=np.uint8), 1.2)
        elif kind == "noise":
            base = rng.integers(0, 256, (H, W + D + 8), dtype=np.uint8)
        elif kind == "binary":
            base = (rng.integers(0, 2, (H, W + D + 8), dtype=np.uint8) * 255).astype(np.uint8)
        else:
            raise ValueError(kind)
        if kind == "smooth":
            d = disparity_field(H, W, D)
            xs = np.arange(W)[None, :] + d
            L = base[:, :W]
            R = np.take_along_axis(base, xs, axis=1)
        else:  # uncorrelated
            L = base[:, :W]
            R = rng.integers(0, 256, (H, W), dtype=np.uint8) if kind == "noise" else \
                (rng.integers(0, 2, (H, W), dtype=np.uint8) * 255).astype(np.uint8)
        outs.append((np.ascontiguousarray(L), np.ascontiguousarray(R)))
    if channels == 1:
        return outs[0]
    return (np.ascontiguousarray(np.stack([o[0] for o in outs], -1)),
            np.ascontiguousarray(np.stack([o[1] for o in outs], -1)))


def sgbm_kwargs(D: int, bs: int = 5, mode: int = 0, channels: int = 1, speckle: bool = True, minD: int = 0) -> dict:
    """Matcher parameters of SURVEY.md §8d."""
    return dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8 * channels * bs * bs,
                P2=32 * channels * bs * bs, disp12MaxDiff=1, preFilterCap=63, uniquenessRatio=10,
                speckleWindowSize=100 if speckle else 0, speckleRange=2 if speckle else 0, mode=mode)


def rodrigues(rv) -> np.ndarray:
    rv = np.asarray(rv, np.float64)
    th = np.linalg.norm(rv)
    if th == 0:
        return np.eye(3)
    k = rv / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx


def synthetic_rig(W: int, H: int, f: float | None = None):
    """Near-parallel rig of SURVEY.md §8d: K1,d1,K2,d2,R,T."""
    if f is None:
        f = 1400.0 if W >= 1920 else (1000.0 if W >= 1280 else 520.0)
    K1 = np.array([[f, 0, W / 2 - 3.2], [0, 1.002 * f, H / 2 + 2.1], [0, 0, 1]])
    K2 = np.array([[1.01 * f, 0, W / 2 + 4.4], [0, 1.008 * f, H / 2 - 1.7], [0, 0, 1]])
    d1 = np.array([-0.12, 0.05, 7e-4, -4e-4, -0.01])
    d2 = np.array([-0.11, 0.04, -5e-4, 6e-4, -0.008])
    R = rodrigues([0.004, -0.012, 0.002])
    T = np.array([-120.0, 0.4, 0.9])
    return K1, d1, K2, d2, R, T


def simple_rectify(K1, K2, R, T, W, H):
    """Host-side stand-in for cv2.stereoRectify (calibration-time work stays on the host; used only to
    make synthetic R1,R2,P1,P2,Q when cv2 is unavailable).  Bouguet-style, horizontal rig."""
    rv = R
    # split rotation in half
    w, V = np.linalg.eig(rv)
    # axis-angle of R
    ang = np.arccos(np.clip((np.trace(rv) - 1) / 2, -1, 1))
    if ang < 1e-12:
        om = np.zeros(3)
    else:
        om = ang / (2 * np.sin(ang)) * np.array([rv[2, 1] - rv[1, 2], rv[0, 2] - rv[2, 0], rv[1, 0] - rv[0, 1]])
    r_r = rodrigues(-0.5 * om)
    t = r_r @ T
    uu = np.array([1.0 if t[0] > 0 else -1.0, 0, 0])
    ww = np.cross(t, uu)
    nw = np.linalg.norm(ww)
    if nw > 0:
        ww *= np.arccos(np.clip(abs(t[0]) / np.linalg.norm(t), -1, 1)) / nw
    wR = rodrigues(ww)
    R1 = wR @ r_r.T
    R2 = wR @ r_r
    tn = R2 @ T
    fnew = min(K1[1, 1], K2[1, 1])
    cx, cy = (W - 1) / 2.0, (H - 1) / 2.0
    P1 = np.array([[fnew, 0, cx, 0], [0, fnew, cy, 0], [0, 0, 1, 0]])
    P2 = P1.copy()
    P2[0, 3] = tn[0] * fnew
    Q = np.array([[1, 0, 0, -cx], [0, 1, 0, -cy], [0, 0, 0, fnew], [0, 0, -1.0 / tn[0], 0]])
    return R1, R2, P1, P2, Q


def boxes_for_frame(H: int, W: int, D: int, n: int = 8, seed: int = 0) -> np.ndarray:
    rng = np.random.default_rng(seed + 77)
    x1 = rng.uniform(D, max(W - 64, D + 1), n)
    y1 = rng.uniform(0, max(H - 64, 1), n)
    w = rng.uniform(32, 256, n)
    h = rng.uniform(32, 256, n)
    b = np.stack([x1, y1, np.minimum(x1 + w, W), np.minimum(y1 + h, H)], 1).astype(np.float32)
    return b
