"""undistortPoints / triangulatePoints / BGR2GRAY / undistort — oracle and CUDA path against
cv2 4.13.0 golden vectors (tests/golden/make_golden_sparse.py).  Tolerances: points are f64 arithmetic rounded to the
input dtype — 1e-5 relative (north_star's float bar; observed <= 1 ulp); integer images bit-exact."""
import os

import numpy as np
import pytest

import make_golden_sparse as mg
import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, "tests", "golden", "golden_sparse.npz"))
C = mg.cases()
DISTS = (("d5", C["d1"]), ("d8", C["d8"]), ("d12", C["d12"]), ("d0", np.zeros(5)))
RTOL = 1e-5


def _close(a, b, scale=None):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    s = np.maximum(np.abs(b), 1.0) if scale is None else scale
    return float(np.max(np.abs(a - b) / s))


def _tri_ok(h, ref):
    h, ref = np.asarray(h, np.float64), np.asarray(ref, np.float64)
    assert h.shape == ref.shape
    assert _close(np.abs(h), np.abs(ref)) < 1e-5                       # unit vector up to sign
    assert _close((h[:3] / h[3]).T, (ref[:3] / ref[3]).T) < RTOL       # what callers use


# ------------------------------------------------------------------ oracle (CPU)
def test_oracle_undistort_points():
    for name, d in DISTS:
        assert _close(oracle.undistort_points(C["pts32"], C["K1"], d, P=C["K1"]), G[f"und32_{name}_PK"].reshape(-1, 2)) < RTOL
        assert _close(oracle.undistort_points(C["pts64"], C["K1"], d), G[f"und64_{name}_none"].reshape(-1, 2)) < 1e-12
        assert _close(oracle.undistort_points(C["pts32"], C["K1"], d, R=C["Rr"], P=C["P34"]),
                      G[f"und32_{name}_RP"].reshape(-1, 2)) < RTOL


def test_oracle_triangulate_and_gray():
    _tri_ok(oracle.triangulate_points(C["P1"], C["P2"], C["a64"].T, C["b64"].T), G["tri64"])
    _tri_ok(oracle.triangulate_points(C["P1"], C["P2"], C["a32"].T, C["b32"].T), G["tri32"])
    assert np.array_equal(oracle.bgr2gray(C["bgr"]), G["gray"])
    assert np.array_equal(oracle.bgr2gray(C["bgr_small"]), G["gray_small"])


# ------------------------------------------------------------------ CUDA path
@pytest.mark.gpu
def test_gpu_undistort_points(sv):
    for name, d in DISTS:
        got = sv.undistortPoints(C["pts32"], C["K1"], d, P=C["K1"])
        assert got.dtype == np.float32 and got.shape == (257, 1, 2)
        assert _close(got, G[f"und32_{name}_PK"]) < RTOL
        got = sv.undistortPoints(C["pts64"], C["K1"], d)
        assert got.dtype == np.float64 and got.shape == (64, 1, 2)   # cv2 returns Nx1x2 for 1xNx2 input too
        assert _close(got, G[f"und64_{name}_none"]) < 1e-12
        assert _close(got.reshape(-1, 2), oracle.undistort_points(C["pts64"], C["K1"], d)) < 1e-13
        assert _close(sv.undistortPoints(C["pts32"], C["K1"], d, R=C["Rr"], P=C["P34"]), G[f"und32_{name}_RP"]) < RTOL
    with pytest.raises(ValueError):
        sv.undistortPoints(C["pts32"].astype(np.int32), C["K1"], C["d1"])


@pytest.mark.gpu
def test_gpu_triangulate_points(sv):
    h = sv.triangulatePoints(C["P1"], C["P2"], C["a64"], C["b64"])
    assert h.dtype == np.float64
    _tri_ok(h, G["tri64"])
    h = sv.triangulatePoints(C["P1"], C["P2"], C["a32"], C["b32"])
    assert h.dtype == np.float32
    _tri_ok(h, G["tri32"])
    # the shapes undistortPoints returns, and a single (2,) point as in ball_drop.py:254
    h = sv.triangulatePoints(C["P1"], C["P2"], C["a64"].T.reshape(-1, 1, 2), C["b64"].T.reshape(-1, 1, 2))
    _tri_ok(h, G["tri64"])
    h1 = sv.triangulatePoints(C["P1"], C["P2"], C["a64"][:, 0], C["b64"][:, 0])
    _tri_ok(h1, G["tri64"][:, :1])


@pytest.mark.gpu
def test_gpu_calculate_3d_position_batched(sv):
    """Batched form of the reference's calculate_3d_position vs cv2 run point by point (ball_drop.py:225-259)."""
    got = sv.triangulate_stereo(C["lp"], C["rp"], C["K1"], C["d1"], C["K2"], C["d2"], C["R"], C["T"])
    assert got.shape == G["calc3d"].shape
    assert _close(got, G["calc3d"], scale=np.linalg.norm(G["calc3d"], axis=1, keepdims=True)) < 2e-5


@pytest.mark.gpu
def test_gpu_bgr2gray_and_undistort(sv):
    import torch
    assert np.array_equal(sv.cvtColor(C["bgr"], sv.COLOR_BGR2GRAY), G["gray"])
    assert np.array_equal(sv.cvtColor(C["bgr_small"], sv.COLOR_BGR2GRAY), G["gray_small"])
    t = torch.from_numpy(np.stack([C["bgr_small"]] * 3)).cuda()
    out = sv.cvtColor(t, sv.COLOR_BGR2GRAY)
    assert out.is_cuda and all(np.array_equal(out[i].cpu().numpy(), G["gray_small"]) for i in range(3))
    Ks = C["K1"] * [[0.1, 1, 0.1], [1, 0.1, 0.1], [1, 1, 1]]
    assert np.array_equal(sv.undistort(C["bgr_small"], Ks, C["d1"]), G["undist_img"])
    assert np.array_equal(sv.undistort(C["bgr_small"][..., 1].copy(), Ks, C["d1"]), G["undist_img"][..., 1])
    m1, m2 = sv.initUndistortRectifyMap(Ks, C["d1"], None, Ks, (200, 120), sv.CV_16SC2)
    assert m1.dtype == np.int16 and m2.dtype == np.uint16
    assert np.array_equal(m1, G["fix_map1"]) and np.array_equal(m2, G["fix_map2"])
    assert np.array_equal(sv.remap(C["bgr_small"], m1, m2, sv.INTER_LINEAR), G["undist_img"])


# ------------------------------------------------------------------ 14-coefficient (tilt) undistortPoints
def _tilt_case():
    import make_golden_tilt as mt
    g = np.load(os.path.join(ROOT, "tests", "golden", "golden_tilt.npz"))
    return mt, g, mt.points()


def test_oracle_undistort_points_with_tilt():
    mt, g, pts = _tilt_case()
    assert _close(oracle.undistort_points(pts, mt.K, mt.DIST), g["plain"].reshape(-1, 2)) < 1e-12
    assert _close(oracle.undistort_points(pts, mt.K, mt.DIST, R=g["R"], P=mt.K), g["rect"].reshape(-1, 2)) < 1e-12


@pytest.mark.gpu
def test_gpu_undistort_points_with_tilt(sv):
    mt, g, pts = _tilt_case()
    assert _close(sv.undistortPoints(pts.reshape(-1, 1, 2), mt.K, mt.DIST), g["plain"]) < 1e-12
    assert _close(sv.undistortPoints(pts.reshape(-1, 1, 2), mt.K, mt.DIST, R=g["R"], P=mt.K), g["rect"]) < 1e-12
    assert _close(sv.undistortPoints(pts.astype(np.float32).reshape(-1, 1, 2), mt.K, mt.DIST, R=g["R"], P=mt.K), g["f32"]) < RTOL
