"""GPU parity for the ops either side of SGBM: maps, remap, median3, filterSpeckles, reproject, box median."""
import hashlib

import numpy as np
import pytest

import cases
import oracle
from stereo_vision_b200 import synth

pytestmark = pytest.mark.gpu


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _q32(a):
    return np.rint(a * np.float32(32)).astype(np.int32)


def test_remap_matches_cv2_golden_bit_exact(sv, golden):
    for i, c in enumerate(cases.remap_cases()):
        src, mx, my = cases.remap_inputs(c)
        got = sv.remap(src, mx, my, sv.INTER_LINEAR)
        assert got.dtype == np.uint8 and got.shape == golden[f"remap_{i}"].shape
        assert np.array_equal(got, golden[f"remap_{i}"]), f"remap case {i}"
        assert np.array_equal(got, oracle.remap(src, mx, my))


def test_remap_1080p_bgr_and_batched_tensor(sv):
    import torch
    W, H = 1920, 1080
    rng = np.random.default_rng(11)
    src = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    K1, d1, K2, d2, R, T = synth.synthetic_rig(W, H)
    R1, R2, P1, P2, Q = synth.simple_rectify(K1, K2, R, T, W, H)
    mx, my = sv.initUndistortRectifyMap(K1, d1, R1, P1, (W, H), sv.CV_32FC1)
    ox, oy = oracle.init_undistort_rectify_map(K1, d1, R1, P1, (W, H))
    assert np.array_equal(_q32(mx), _q32(ox)) and np.array_equal(_q32(my), _q32(oy))
    got = sv.remap(src, mx, my, sv.INTER_LINEAR)
    assert np.array_equal(got, oracle.remap(src, ox, oy))
    # identity map returns the source
    ix, iy = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    assert np.array_equal(sv.remap(src, ix, iy, sv.INTER_LINEAR), src)
    # batched torch tensors, zero-copy, odd width (scalar tail path)
    g = rng.integers(0, 256, (3, 97, 131), dtype=np.uint8)
    mxs = (np.arange(131)[None, :] + rng.uniform(-3, 3, (97, 131))).astype(np.float32)
    mys = (np.arange(97)[:, None] + rng.uniform(-3, 3, (97, 131))).astype(np.float32)
    out = sv.remap(torch.from_numpy(g).cuda(), torch.from_numpy(mxs).cuda(), torch.from_numpy(mys).cuda())
    assert out.is_cuda
    for b in range(3):
        assert np.array_equal(out[b].cpu().numpy(), oracle.remap(g[b], mxs, mys))


def test_maps_match_cv2_golden(sv, golden, golden_full):
    for (W, H) in [(640, 480), (1280, 720), (1920, 1080)]:
        K1, d1, K2, d2, R, T = synth.synthetic_rig(W, H)
        rect = golden[f"rect_{W}"]
        R1, R2 = rect[0:9].reshape(3, 3), rect[9:18].reshape(3, 3)
        P1, P2 = rect[18:30].reshape(3, 4), rect[30:42].reshape(3, 4)
        for side, (K, d, Rr, P) in enumerate([(K1, d1, R1, P1), (K2, d2, R2, P2)]):
            mx, my = sv.initUndistortRectifyMap(K, d, Rr, P, (W, H), sv.CV_32FC1)
            assert mx.dtype == np.float32 and mx.shape == (H, W)
            # bit-exact after the x32 quantisation remap applies
            assert [_sha(_q32(mx)), _sha(_q32(my))] == golden_full[f"mapq_{W}_{side}"]
            if W == 640:
                assert (mx[::8, ::8] != golden[f"mapx_{W}_{side}"]).sum() <= 1
                assert (my[::8, ::8] != golden[f"mapy_{W}_{side}"]).sum() <= 1
    # 14-coefficient model incl. tilt, R=None, P=None
    K1, d1, *_ = synth.synthetic_rig(640, 480)
    d14 = np.concatenate([d1, [0.01, -0.02, 0.003, 1e-3, -2e-3, 3e-3, 1e-3, 0.01, -0.02]])
    mx, my = sv.initUndistortRectifyMap(K1, d14, None, None, (640, 480), sv.CV_32FC1)
    ox, oy = oracle.init_undistort_rectify_map(K1, d14, None, None, (640, 480))
    assert np.array_equal(mx, ox) and np.array_equal(my, oy)
    with pytest.raises(ValueError):
        sv.initUndistortRectifyMap(K1, np.zeros(6), None, None, (640, 480), sv.CV_32FC1)


def test_median_and_speckle_match_cv2_golden(sv, golden):
    for j, arr in enumerate(cases.post_inputs()):
        assert np.array_equal(sv.medianBlur(arr, 3), golden[f"median_{j}"])
        for (sz, df) in cases.SPECKLE_PARAMS:
            a = arr.copy()
            r, _ = sv.filterSpeckles(a, -16, sz, df)
            assert r is a and np.array_equal(a, golden[f"speckle_{j}_{sz}_{df}"]), (j, sz, df)


def test_speckle_large_components_1080p(sv):
    rng = np.random.default_rng(3)
    H, W = 1080, 1920
    a = (np.kron(rng.integers(0, 30, (H // 40, W // 40)), np.ones((40, 40))).astype(np.int16) * 16)
    a += rng.integers(-6, 7, (H, W)).astype(np.int16)
    a[rng.random((H, W)) < 0.05] = -16
    a[200:900, 300:1500] = 77 * 16           # one huge component with a snake-like hole pattern
    a[200:900:4, 300:1400] = -16
    ref = oracle.filter_speckles(a, -16, 100, 32)
    got = a.copy()
    sv.filterSpeckles(got, -16, 100, 32)
    assert np.array_equal(got, ref)
    assert np.array_equal(sv.medianBlur(a, 3), oracle.median3(a))


def test_reproject_matches_cv2_golden_within_1e5(sv, golden):
    W, H = 640, 480
    rng = np.random.default_rng(5)
    Q = golden["rect_640"][-16:].reshape(4, 4)
    disp = rng.integers(-16, 64 * 16, (H, W)).astype(np.int16)
    for hm in (0, 1):
        for tag, d in (("s16", disp), ("f32", disp.astype(np.float32) / 16)):
            got = sv.reprojectImageTo3D(d, Q, handleMissingValues=bool(hm))
            assert got.dtype == np.float32 and got.shape == (H, W, 3)
            ref = golden[f"reproj_{tag}_{hm}"]
            sub = got[::4, ::4]
            fin = np.isfinite(ref)
            assert np.array_equal(np.isfinite(sub), fin)
            rel = np.abs(sub[fin] - ref[fin]) / np.maximum(np.abs(ref[fin]), 1e-30)
            assert rel.max() <= 1e-5          # north_star tolerance for reprojected XYZ
            assert np.array_equal(got, oracle.reproject_to_3d(d, Q, bool(hm)), equal_nan=True)
    for dt in (np.uint8, np.int32):
        d = rng.integers(0, 200, (33, 47)).astype(dt)
        assert np.array_equal(sv.reprojectImageTo3D(d, Q), oracle.reproject_to_3d(d, Q), equal_nan=True)
    with pytest.raises(ValueError):
        sv.reprojectImageTo3D(disp.astype(np.float64), Q)


def test_box_median_depth(sv):
    import torch
    H, W, D = 480, 640, 64
    L, R = synth.stereo_pair(H, W, D, seed=1000)
    kw = synth.sgbm_kwargs(D, 5, 0)
    disp = oracle.sgbm_compute(L, R, **kw)
    K1, d1, K2, d2, Rm, T = synth.synthetic_rig(W, H)
    Q = synth.simple_rectify(K1, K2, Rm, T, W, H)[4]
    boxes = synth.boxes_for_frame(H, W, D, 8, seed=1)
    extra = np.array([[0, 0, 10, 10], [-50, -50, 700, 500], [100.2, 50.7, 100.9, 50.9], [630, 470, 660, 490],
                      [300, 200, 300, 260]], np.float32)   # all-invalid band, whole image, 1 px, clipped, empty
    boxes = np.concatenate([boxes, extra])
    ref = oracle.box_median_depth(disp, Q, boxes, -16)
    got = sv.box_median_depth(disp, Q, boxes, minDisparity=0)
    assert np.array_equal(got, ref, equal_nan=True)
    assert np.isnan(got[8]) and np.isnan(got[-1])
    # consistent with reprojectImageTo3D + np.median
    xyz = sv.reprojectImageTo3D(disp.astype(np.float32) / 16, Q)
    x1, y1, x2, y2 = [int(v) for v in (np.floor(boxes[0, 0]), np.floor(boxes[0, 1]), np.ceil(boxes[0, 2]), np.ceil(boxes[0, 3]))]
    z = xyz[y1:y2, x1:x2, 2][(disp[y1:y2, x1:x2] != -16) & np.isfinite(xyz[y1:y2, x1:x2, 2])]
    assert np.float32(np.median(z)) == got[0]
    # batched maps with per-box frame index
    db = torch.from_numpy(np.stack([disp, disp[::-1].copy()])).cuda()
    bi = np.array([0, 1, 0, 1], np.int32)
    gb = sv.box_median_depth(db, Q, boxes[:4], box_image=bi)
    assert gb.is_cuda
    r1 = oracle.box_median_depth(disp[::-1].copy(), Q, boxes[:4], -16)
    assert np.array_equal(gb.cpu().numpy(), np.where(bi == 0, ref[:4], r1), equal_nan=True)


def test_reference_held_calibration_fixture_maps_and_remap(sv):
    """The only numeric fixture the reference holds for this path (calibration_tests/calibrate_v5: negative focal length
    in rect_P1 -> maps far outside the image; 5-coefficient distortion): x32-quantised map hashes and remapped frame
    hashes made by cv2 (tests/golden/make_golden_r2.py)."""
    import hashlib
    import json
    import os
    import cases
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    with open(os.path.join(os.path.dirname(__file__), "golden", "golden_r2.json")) as f:
        m = json.load(f)["ref_v5"]
    fx = {k: np.array(v) for k, v in m["fixture"].items()}
    W, H = m["size"]
    q = lambda a: np.rint(a * np.float32(32)).astype(np.int32)
    src = cases.ref_v5_frame()
    for cam in (1, 2):
        for tag, P in (("", fx[f"P{cam}"]), ("_K", fx[f"K{cam}"])):
            mx, my = sv.initUndistortRectifyMap(fx[f"K{cam}"], fx[f"d{cam}"], fx[f"R{cam}"], P, (W, H), sv.CV_32FC1)
            assert [sha(q(mx)), sha(q(my))] == m[f"mapq{tag}_{cam}"], (cam, tag)
            assert sha(sv.remap(src, mx, my, sv.INTER_LINEAR)) == m[f"remap{tag}_{cam}"], (cam, tag)


def test_chain_c2_pipeline_matches_cv2_link_by_link(sv):
    """c2 end to end on the device: raw 1280x720 BGR -> remap x2 -> BGR2GRAY x2 -> StereoSGBM.compute ->
    reprojectImageTo3D -> box medians; rectified / gray images and disparity by cv2 hash, XYZ and medians <= 1e-5."""
    import hashlib
    import json
    import os
    import torch
    import cases
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    gd = os.path.join(os.path.dirname(__file__), "golden")
    with open(os.path.join(gd, "golden_r2.json")) as f:
        c = json.load(f)["chain_c2"]
    npz = np.load(os.path.join(gd, "golden_r2.npz"))
    W, H = 1280, 720
    L, R = cases.chain_c2_inputs()
    assert [sha(L), sha(R)] == c["in_sha"]
    K1, d1, K2, d2, _, _ = synth.synthetic_rig(W, H)
    rc = {k: np.array(v) for k, v in c["rectify"].items()}
    dev = "cuda:0"
    m1 = sv.initUndistortRectifyMap(K1, d1, rc["R1"], rc["P1"], (W, H), sv.CV_32FC1, device=dev)
    m2 = sv.initUndistortRectifyMap(K2, d2, rc["R2"], rc["P2"], (W, H), sv.CV_32FC1, device=dev)
    Ld, Rd = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
    Lr, Rr = sv.remap(Ld, *m1, sv.INTER_LINEAR), sv.remap(Rd, *m2, sv.INTER_LINEAR)      # everything stays on the device
    assert [sha(Lr.cpu().numpy()), sha(Rr.cpu().numpy())] == c["rect"]
    Lg, Rg = sv.cvtColor(Lr, sv.COLOR_BGR2GRAY), sv.cvtColor(Rr, sv.COLOR_BGR2GRAY)
    assert [sha(Lg.cpu().numpy()), sha(Rg.cpu().numpy())] == c["gray"]
    disp = sv.StereoSGBM_create(**c["kw"]).compute(Lg, Rg)
    assert sha(disp.cpu().numpy()) == c["disp"]
    xyz = sv.reprojectImageTo3D(disp.to(torch.float32) / 16, rc["Q"]).cpu().numpy()[::8, ::8]
    ref = npz["chain_c2_xyz"]
    fin = np.isfinite(ref)
    assert np.array_equal(np.isfinite(xyz), fin)
    assert (np.abs(xyz[fin] - ref[fin]) / np.maximum(np.abs(ref[fin]), 1e-30)).max() <= 1e-5
    med = sv.box_median_depth(disp, rc["Q"], npz["chain_c2_boxes"]).cpu().numpy()
    refm = npz["chain_c2_med"]
    assert np.array_equal(np.isnan(med), np.isnan(refm))
    ok = ~np.isnan(refm)
    assert (np.abs(med[ok] - refm[ok]) / np.abs(refm[ok])).max() <= 1e-5
