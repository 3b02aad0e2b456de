"""Calibration loader and rectify glue: reads the reference's `results/` layout, takes the
calibration-time scalars (R1, R2, P1, P2, Q) from the host and keeps everything per-frame on the device.

Reference interfaces mirrored here:
  load_calibration(test_dir)           stereo_calibration/scripts/ball_drop.py:162-223 (same files, same fallbacks,
                                       same keys, `None` + a printed error when something is missing)
  StereoRig.from_calibration(...)      checkerboard_distance.py:49-57 (stereoRectify alpha=0.0 -> two CV_32FC1 map pairs)
  StereoRig.rectify / depth / boxes    checkerboard_distance.py:71-72 (remap), ball_drop.py:225-259 (3-D per detection)

`cv2.stereoRectify` is calibration-time host work (north_star: the calibration scripts stay on the host unchanged):
pass its outputs in, or hand `from_calibration` the host function (`stereo_rectify=cv2.stereoRectify`).  Nothing here runs per frame
on the CPU.
"""
from __future__ import annotations

import os
import pickle

import numpy as np
import torch

from . import api

CALIB_ZERO_DISPARITY = 1024


def load_calibration(test_dir):
    """Intrinsics + extrinsics of one test directory -> dict(left_matrix, left_dist, right_matrix, right_dist, R, T)
    or None (with the reference's error line printed) — ball_drop.py:162-223."""
    results_dir = os.path.join(test_dir, "results")
    intrinsic_dir = os.path.join(results_dir, "intrinsic_params")
    extrinsic_dir = os.path.join(results_dir, "extrinsic_params")
    if not os.path.exists(intrinsic_dir) or not os.path.exists(extrinsic_dir):
        print(f"Error: Calibration directories not found at {results_dir}")
        return None
    try:
        lp, rp = (os.path.join(intrinsic_dir, f"{s}_intrinsics.pkl") for s in ("left", "right"))
        if os.path.exists(lp) and os.path.exists(rp):
            with open(lp, "rb") as f:
                left_matrix, left_dist = pickle.load(f)
            with open(rp, "rb") as f:
                right_matrix, right_dist = pickle.load(f)
        else:
            left_matrix = np.loadtxt(os.path.join(intrinsic_dir, "left_matrix.txt"))
            left_dist = np.loadtxt(os.path.join(intrinsic_dir, "left_distortion.txt"))
            right_matrix = np.loadtxt(os.path.join(intrinsic_dir, "right_matrix.txt"))
            right_dist = np.loadtxt(os.path.join(intrinsic_dir, "right_distortion.txt"))
    except Exception as e:  # the reference reports and returns None for any failure here
        print(f"Error loading intrinsic parameters: {e}")
        return None
    try:
        ep = os.path.join(extrinsic_dir, "extrinsic_params.pkl")
        if os.path.exists(ep):
            with open(ep, "rb") as f:
                d = pickle.load(f)
            R, T = d["R"], d["T"]
        else:
            R = np.loadtxt(os.path.join(extrinsic_dir, "stereo_rotation_matrix.txt"))
            T = np.loadtxt(os.path.join(extrinsic_dir, "stereo_translation_vector.txt"))
            if T.shape == (3,):
                T = T.reshape(3, 1)
    except Exception as e:
        print(f"Error loading extrinsic parameters: {e}")
        return None
    return {"left_matrix": left_matrix, "left_dist": left_dist, "right_matrix": right_matrix,
            "right_dist": right_dist, "R": R, "T": T}


def _host_stereo_rectify(stereo_rectify, calib, image_size, alpha, flags):
    """One call of the host's stereoRectify (e.g. `cv2.stereoRectify`, handed in by the caller: this package never
    imports the CPU library) with the reference's argument order (checkerboard_distance.py:49-51)."""
    out = stereo_rectify(
        np.asarray(calib["left_matrix"], np.float64), np.asarray(calib["left_dist"], np.float64),
        np.asarray(calib["right_matrix"], np.float64), np.asarray(calib["right_dist"], np.float64),
        tuple(int(v) for v in image_size), np.asarray(calib["R"], np.float64),
        np.asarray(calib["T"], np.float64).reshape(3, 1), flags=flags, alpha=alpha)
    return tuple(out[:5])


class StereoRig:
    """A calibrated, rectified stereo pair resident on one GPU: both map pairs, Q and an SGBM matcher.
    Frames go in as uint8 HxW / HxWx3 (NumPy or CUDA tensors, or batches [B,H,W(,3)] of tensors)."""

    def __init__(self, calib, image_size, rectification, device=None, matcher=None):
        self.calib = {k: np.asarray(v, np.float64) for k, v in calib.items()}
        self.image_size = (int(image_size[0]), int(image_size[1]))
        self.R1, self.R2, self.P1, self.P2, self.Q = (np.asarray(m, np.float64) for m in rectification)
        if self.R1.shape != (3, 3) or self.R2.shape != (3, 3) or self.P1.shape != (3, 4) or self.P2.shape != (3, 4) \
                or self.Q.shape != (4, 4):
            raise ValueError("rectification must be (R1 3x3, R2 3x3, P1 3x4, P2 3x4, Q 4x4)")
        dev = torch.cuda.current_device() if device is None else (torch.device(device).index or 0)
        self.device = dev
        d = f"cuda:{dev}"
        c = self.calib
        self.map_left = api.initUndistortRectifyMap(c["left_matrix"], c["left_dist"], self.R1, self.P1, self.image_size,
                                                    api.CV_32FC1, device=d)
        self.map_right = api.initUndistortRectifyMap(c["right_matrix"], c["right_dist"], self.R2, self.P2,
                                                     self.image_size, api.CV_32FC1, device=d)
        self.matcher = matcher

    @classmethod
    def from_calibration(cls, calib, image_size, alpha=0.0, flags=CALIB_ZERO_DISPARITY, rectification=None,
                         stereo_rectify=None, device=None, matcher=None):
        """image_size = (width, height) as for cv2.  Give either `rectification` = (R1, R2, P1, P2, Q) as returned by
        cv2.stereoRectify, or `stereo_rectify=cv2.stereoRectify` to have it called once here with `alpha`/`flags`
        (checkerboard_distance.py:49-51: alpha=0.0)."""
        if calib is None:
            raise ValueError("no calibration (load_calibration returned None)")
        if rectification is None:
            if stereo_rectify is None:
                raise ValueError("stereoRectify is calibration-time host work: pass rectification=(R1, R2, P1, P2, Q) "
                                 "or stereo_rectify=cv2.stereoRectify")
            rectification = _host_stereo_rectify(stereo_rectify, calib, image_size, alpha, flags)
        return cls(calib, image_size, rectification, device=device, matcher=matcher)

    @classmethod
    def from_dir(cls, test_dir, image_size, **kw):
        return cls.from_calibration(load_calibration(test_dir), image_size, **kw)

    # ---- per-frame work: all on the device ----
    def _dev(self, img):
        return img if isinstance(img, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(img)).to(f"cuda:{self.device}")

    def rectify(self, left, right):
        """remap both views with the rig's maps (checkerboard_distance.py:71-72).  Returns CUDA tensors."""
        with torch.cuda.device(self.device):
            return (api.remap(self._dev(left), *self.map_left, api.INTER_LINEAR),
                    api.remap(self._dev(right), *self.map_right, api.INTER_LINEAR))

    def disparity(self, left, right, rectified=False):
        """rectify -> (BGR->gray when 3-channel frames should be matched as gray: pass gray frames for that) -> SGBM."""
        if self.matcher is None:
            raise ValueError("StereoRig was built without a matcher (pass matcher=StereoSGBM_create(...))")
        L, R = (self._dev(left), self._dev(right)) if rectified else self.rectify(left, right)
        batched = L.dim() == 4 or (L.dim() == 3 and L.shape[-1] not in (1, 3))
        with torch.cuda.device(self.device):
            return self.matcher.compute_batch(L, R) if batched else self.matcher.compute(L, R)

    def box_depth(self, left, right, boxes, rectified=False, box_image=None):
        """Median depth Z of each detector box of the LEFT rectified view: the dense replacement for the reference's
        centre-point triangulation.  For a batch of frames `box_image` (int32, one frame index per box) is required."""
        disp = self.disparity(left, right, rectified)
        if disp.dim() == 3 and box_image is None:
            raise ValueError("box_depth on a batch of frames needs box_image (the frame index of every box)")
        return api.box_median_depth(disp, self.Q, boxes, minDisparity=self.matcher.getMinDisparity(), box_image=box_image)

    def triangulate(self, left_pts, right_pts, rectified_frame=False):
        """Batched 3-D positions of matched detections, (N,2) pixel coordinates in the ORIGINAL images.
        rectified_frame=False: calculate_3d_position (ball_drop.py:225-259) — camera-1 frame, P = K[I|0], K[R|T].
        rectified_frame=True : checkerboard_distance.py:83-103 — undistortPoints(R=R1,P=P1) + triangulate with P1/P2."""
        c = self.calib
        with torch.cuda.device(self.device):
            if not rectified_frame:
                return api.triangulate_stereo(left_pts, right_pts, c["left_matrix"], c["left_dist"], c["right_matrix"],
                                              c["right_dist"], c["R"], c["T"])
            lp = np.asarray(left_pts, np.float32).reshape(-1, 1, 2) if not isinstance(left_pts, torch.Tensor) else left_pts
            rp = np.asarray(right_pts, np.float32).reshape(-1, 1, 2) if not isinstance(right_pts, torch.Tensor) else right_pts
            lu = api.undistortPoints(lp, c["left_matrix"], c["left_dist"], R=self.R1, P=self.P1)
            ru = api.undistortPoints(rp, c["right_matrix"], c["right_dist"], R=self.R2, P=self.P2)
            X = api.triangulatePoints(self.P1, self.P2, lu, ru)
            return (X[:3] / X[3]).T
