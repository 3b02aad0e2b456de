"""CPU: pins the oracle (oracle/sv_oracle.c) against the committed cv2 4.13.0 golden vectors."""
import hashlib

import numpy as np
import pytest

import cases
import oracle
from stereo_vision_b200 import synth


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_sgbm_small_sweep(golden):
    for i, c in enumerate(cases.sgbm_small_cases()):
        L, R = cases.sgbm_inputs(c)
        key = np.frombuffer(bytes.fromhex(_sha(L)[:16] + _sha(R)[:16]), np.uint8)
        assert np.array_equal(key, golden[f"sgbm_in_{i}"]), "seeded inputs not reproducible"
        assert np.array_equal(oracle.sgbm_compute(L, R, **c["kw"]), golden[f"sgbm_{i}"]), (i, c)


@pytest.mark.parametrize("name", ["c1", "c2"])
def test_sgbm_full_size_hash(golden_full, name):
    L, R, kw = cases.full_inputs(name)
    assert [_sha(L), _sha(R)] == golden_full["full"][name]["in_sha256"]
    assert _sha(oracle.sgbm_compute(L, R, **kw)) == golden_full["full"][name]["sha256"]


def test_remap(golden):
    for i, c in enumerate(cases.remap_cases()):
        src, mx, my = cases.remap_inputs(c)
        assert np.array_equal(oracle.remap(src, mx, my), golden[f"remap_{i}"])


def test_maps(golden, golden_full):
    q = lambda a: np.rint(a * np.float32(32)).astype(np.int32)
    for (W, H) in [(640, 480), (1280, 720)]:
        K1, d1, K2, d2, R, T = synth.synthetic_rig(W, H)
        rect = golden[f"rect_{W}"]
        R1, R2 = rect[0:9].reshape(3, 3), rect[9:18].reshape(3, 3)
        P1, P2 = rect[18:30].reshape(3, 4), rect[30:42].reshape(3, 4)
        for side, (K, d, Rr, P) in enumerate([(K1, d1, R1, P1), (K2, d2, R2, P2)]):
            mx, my = oracle.init_undistort_rectify_map(K, d, Rr, P, (W, H))
            assert [_sha(q(mx)), _sha(q(my))] == golden_full[f"mapq_{W}_{side}"]


def test_post_filters(golden):
    for j, arr in enumerate(cases.post_inputs()):
        assert np.array_equal(oracle.median3(arr), golden[f"median_{j}"])
        for (sz, df) in cases.SPECKLE_PARAMS:
            assert np.array_equal(oracle.filter_speckles(arr, -16, sz, df), golden[f"speckle_{j}_{sz}_{df}"])


def test_reproject(golden):
    rng = np.random.default_rng(5)
    Q = golden["rect_640"][-16:].reshape(4, 4)
    disp = rng.integers(-16, 64 * 16, (480, 640)).astype(np.int16)
    for hm in (0, 1):
        for tag, d in (("s16", disp), ("f32", disp.astype(np.float32) / 16)):
            got = oracle.reproject_to_3d(d, Q, bool(hm))[::4, ::4]
            ref = golden[f"reproj_{tag}_{hm}"]
            fin = np.isfinite(ref)
            assert np.array_equal(np.isfinite(got), fin)
            assert (np.abs(got[fin] - ref[fin]) / np.maximum(np.abs(ref[fin]), 1e-30)).max() <= 1e-5


def test_box_median_semantics():
    disp = np.full((20, 30), -16, np.int16)
    disp[5:10, 5:10] = np.arange(25).reshape(5, 5) * 16 + 160
    Q = np.array([[1, 0, 0, -15], [0, 1, 0, -10], [0, 0, 0, 500.0], [0, 0, 1 / 100.0, 0]])
    out = oracle.box_median_depth(disp, Q, np.array([[5, 5, 10, 10], [0, 0, 4, 4], [4.5, 4.5, 6.2, 6.2]], np.float32), -16)
    z = 500.0 / ((np.arange(25) + 10) / 100.0)
    assert out[0] == np.float32(np.median(z.astype(np.float32)))
    assert np.isnan(out[1])
    sub = (disp[4:7, 4:7][disp[4:7, 4:7] != -16] / 16.0)
    assert out[2] == np.float32(np.median((500.0 / (sub / 100.0)).astype(np.float32)))


def test_sgbm_hh4_small_sweep_and_full_hash():
    """MODE_HH4 (mode 3): 4 paths, bottom blockSize//2 rows of the cost volume stay 0."""
    import json
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_hh4.npz"))
    for i, c in enumerate(cases.sgbm_hh4_cases()):
        L, R = cases.sgbm_inputs(c)
        assert np.array_equal(oracle.sgbm_compute(L, R, **c["kw"]), g[f"hh4_{i}"]), (i, c)
    with open(os.path.join(os.path.dirname(__file__), "golden", "golden_hh4_full.json")) as f:
        full = json.load(f)
    L, R, kw = cases.full_inputs("c3")
    assert _sha(oracle.sgbm_compute(L, R, **dict(kw, mode=3))) == full["c3_hh4"]


def _gdir():
    import os
    return os.path.join(os.path.dirname(__file__), "golden")


def test_sgbm_3way_small_sweep_and_full_hashes():
    """MODE_SGBM_3WAY (mode 2): 4 row stripes with restarted paths, 3 paths, the SIMD tie rule and
    the integer uniqueness threshold of cv2's 3-way selection (make_golden_3way.py: thread-count independent)."""
    import json
    import os
    g = np.load(os.path.join(_gdir(), "golden_3way.npz"))
    for i, c in enumerate(cases.sgbm_3way_cases()):
        L, R = cases.sgbm_inputs(c)
        assert np.array_equal(oracle.sgbm_compute(L, R, **c["kw"]), g[f"w3_{i}"]), (i, c)
    with open(os.path.join(_gdir(), "golden_3way_full.json")) as f:
        full = json.load(f)
    for name in ("c1", "c2"):
        L, R, kw = cases.full_inputs(name)
        assert _sha(oracle.sgbm_compute(L, R, **dict(kw, mode=2))) == full[f"{name}_3way"]
    # images too short for the stripe overlap are outside the contract (cv2 behaves differently there)
    z = np.zeros((12, 80), np.uint8)
    with pytest.raises(ValueError):
        oracle.sgbm_compute(z, z, numDisparities=16, blockSize=9, mode=2)


@pytest.fixture(scope="module")
def golden_r2():
    import json
    import os
    with open(os.path.join(_gdir(), "golden_r2.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(_gdir(), "golden_r2.npz"))


def test_reference_held_calibration_fixture(golden_r2):
    """Maps + remap on the numbers the reference itself holds (calibration_tests/calibrate_v5: negative focal length
    in rect_P1, 5-coefficient distortion) — cv2 hashes made by make_golden_r2.py."""
    meta, _ = golden_r2
    m = meta["ref_v5"]
    fx = {k: np.array(v) for k, v in m["fixture"].items()}
    assert fx["P1"][0, 0] < 0                       # the fixture's own oddity: negative focal length
    W, H = m["size"]
    q = lambda a: np.rint(a * np.float32(32)).astype(np.int32)
    src = cases.ref_v5_frame()
    for cam in (1, 2):
        for tag, P in (("", fx[f"P{cam}"]), ("_K", fx[f"K{cam}"])):
            mx, my = oracle.init_undistort_rectify_map(fx[f"K{cam}"], fx[f"d{cam}"], fx[f"R{cam}"], P, (W, H))
            assert [_sha(q(mx)), _sha(q(my))] == m[f"mapq{tag}_{cam}"], (cam, tag)
            assert _sha(oracle.remap(src, mx, my)) == m[f"remap{tag}_{cam}"], (cam, tag)


def test_chain_c2_pipeline(golden_r2):
    """Raw 1280x720 BGR pair -> remap x2 -> BGR2GRAY x2 -> SGBM -> reprojectImageTo3D -> box medians, link by link
    against cv2 (hashes), XYZ within 1e-5 relative."""
    meta, npz = golden_r2
    c = meta["chain_c2"]
    W, H = 1280, 720
    L, R = cases.chain_c2_inputs()
    assert [_sha(L), _sha(R)] == c["in_sha"]
    K1, d1, K2, d2, _, _ = synth.synthetic_rig(W, H)
    rc = {k: np.array(v) for k, v in c["rectify"].items()}
    m1 = oracle.init_undistort_rectify_map(K1, d1, rc["R1"], rc["P1"], (W, H))
    m2 = oracle.init_undistort_rectify_map(K2, d2, rc["R2"], rc["P2"], (W, H))
    Lr, Rr = oracle.remap(L, *m1), oracle.remap(R, *m2)
    assert [_sha(Lr), _sha(Rr)] == c["rect"]
    Lg, Rg = oracle.bgr2gray(Lr), oracle.bgr2gray(Rr)
    assert [_sha(Lg), _sha(Rg)] == c["gray"]
    disp = oracle.sgbm_compute(Lg, Rg, **c["kw"])
    assert _sha(disp) == c["disp"]
    xyz = oracle.reproject_to_3d(disp.astype(np.float32) / 16, rc["Q"])[::8, ::8]
    ref = npz["chain_c2_xyz"]
    fin = np.isfinite(ref)
    assert np.array_equal(np.isfinite(xyz), fin)
    assert (np.abs(xyz[fin] - ref[fin]) / np.maximum(np.abs(ref[fin]), 1e-30)).max() <= 1e-5
    med = oracle.box_median_depth(disp, rc["Q"], npz["chain_c2_boxes"], -16)
    refm = npz["chain_c2_med"]
    assert np.array_equal(np.isnan(med), np.isnan(refm))
    ok = ~np.isnan(refm)
    assert (np.abs(med[ok] - refm[ok]) / np.abs(refm[ok])).max() <= 1e-5
