"""Calibration loader + rectify glue — the reference's results/ layout (ball_drop.py:162-223) and
the checkerboard_distance.py rectify/triangulate sequence against cv2 4.13.0 golden vectors."""
import os
import pickle

import numpy as np
import pytest

import make_golden_sparse as mg
from stereo_vision_b200 import calib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, "tests", "golden", "golden_sparse.npz"))
C = mg.cases()
RIG = C["rig"]


def _write_txt(root, rig, t_flat=True):
    i, e = os.path.join(root, "results", "intrinsic_params"), os.path.join(root, "results", "extrinsic_params")
    os.makedirs(i), os.makedirs(e)
    np.savetxt(os.path.join(i, "left_matrix.txt"), rig["left_matrix"])
    np.savetxt(os.path.join(i, "left_distortion.txt"), rig["left_dist"])
    np.savetxt(os.path.join(i, "right_matrix.txt"), rig["right_matrix"])
    np.savetxt(os.path.join(i, "right_distortion.txt"), rig["right_dist"])
    np.savetxt(os.path.join(e, "stereo_rotation_matrix.txt"), rig["R"])
    np.savetxt(os.path.join(e, "stereo_translation_vector.txt"), rig["T"].ravel() if t_flat else rig["T"])
    return i, e


def test_load_calibration_text_layout(tmp_path):
    _write_txt(str(tmp_path), RIG)
    c = calib.load_calibration(str(tmp_path))
    assert set(c) == {"left_matrix", "left_dist", "right_matrix", "right_dist", "R", "T"}
    for k in c:
        assert np.allclose(c[k], RIG[k], rtol=0, atol=0), k
    assert c["T"].shape == (3, 1)          # flat text vector is reshaped, as the reference does


def test_load_calibration_pickles_win(tmp_path):
    i, e = _write_txt(str(tmp_path), RIG)
    other = {k: v * 1.5 for k, v in RIG.items()}
    with open(os.path.join(i, "left_intrinsics.pkl"), "wb") as f:
        pickle.dump((other["left_matrix"], other["left_dist"]), f)
    c = calib.load_calibration(str(tmp_path))
    assert np.array_equal(c["left_matrix"], RIG["left_matrix"])   # only ONE pickle present: text files are used
    with open(os.path.join(i, "right_intrinsics.pkl"), "wb") as f:
        pickle.dump((other["right_matrix"], other["right_dist"]), f)
    with open(os.path.join(e, "extrinsic_params.pkl"), "wb") as f:
        pickle.dump({"R": other["R"], "T": other["T"]}, f)
    c = calib.load_calibration(str(tmp_path))
    for k in c:
        assert np.array_equal(c[k], other[k]), k


def test_load_calibration_errors(tmp_path, capsys):
    assert calib.load_calibration(str(tmp_path)) is None
    assert "Calibration directories not found" in capsys.readouterr().out
    i, e = _write_txt(str(tmp_path), RIG)
    os.remove(os.path.join(i, "right_distortion.txt"))
    assert calib.load_calibration(str(tmp_path)) is None
    assert "Error loading intrinsic parameters" in capsys.readouterr().out
    np.savetxt(os.path.join(i, "right_distortion.txt"), RIG["right_dist"])
    os.remove(os.path.join(e, "stereo_rotation_matrix.txt"))
    assert calib.load_calibration(str(tmp_path)) is None
    assert "Error loading extrinsic parameters" in capsys.readouterr().out
    with pytest.raises(ValueError):
        calib.StereoRig.from_calibration(None, (320, 180))


@pytest.mark.gpu
def test_gpu_rig_rectify_and_triangulate(sv, tmp_path):
    _write_txt(str(tmp_path), RIG)
    rect = tuple(G[f"rig_{k}"] for k in ("R1", "R2", "P1", "P2", "Q"))
    rig = calib.StereoRig.from_dir(str(tmp_path), (320, 180), rectification=rect,
                                   matcher=sv.StereoSGBM_create(minDisparity=0, numDisparities=32, blockSize=5))
    L, R = rig.rectify(C["rig_left"], C["rig_right"])
    assert L.is_cuda and np.array_equal(L.cpu().numpy(), G["rig_rect_left"])
    assert np.array_equal(R.cpu().numpy(), G["rig_rect_right"])
    # corners -> undistortPoints(R1,P1) -> triangulate with P1/P2 (checkerboard_distance.py:83-103), 1e-5 relative
    X = rig.triangulate(C["rig_lp"], C["rig_rp"], rectified_frame=True)
    ref = G["rig_pts3d"]
    assert np.max(np.abs(X - ref) / np.linalg.norm(ref, axis=1, keepdims=True)) < 2e-5
    # dense path through the same rig: batch == single frames, boxes give finite depth where disparity is valid
    g1 = sv.cvtColor(L, sv.COLOR_BGR2GRAY)
    g2 = sv.cvtColor(R, sv.COLOR_BGR2GRAY)
    d = rig.disparity(g1, g2, rectified=True)
    import torch
    db = rig.disparity(torch.stack([g1, g1]), torch.stack([g2, g2]), rectified=True)
    assert d.shape == (180, 320) and torch.equal(db[0], d) and torch.equal(db[1], d)


@pytest.mark.gpu
def test_gpu_rig_host_stereo_rectify(sv):
    """from_calibration(stereo_rectify=cv2.stereoRectify) calls the host function once (calibration-time)."""
    cv2 = pytest.importorskip("cv2")
    with pytest.raises(ValueError):
        calib.StereoRig.from_calibration(RIG, (320, 180))
    rig = calib.StereoRig.from_calibration(RIG, (320, 180), alpha=0.0, flags=cv2.CALIB_ZERO_DISPARITY,
                                           stereo_rectify=cv2.stereoRectify)
    for k in ("R1", "R2", "P1", "P2", "Q"):
        assert np.allclose(getattr(rig, k), G[f"rig_{k}"], rtol=1e-12, atol=1e-12), k
