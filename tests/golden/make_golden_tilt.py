"""cv2.undistortPoints with all 14 distortion coefficients (tilt included), from the cv2 4.13.0 binary.
Run here:  python tests/golden/make_golden_tilt.py -> tests/golden/golden_tilt.npz"""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
K = np.array([[900., 0, 640], [0, 905, 360], [0, 0, 1]])
DIST = np.array([-0.12, 0.05, 7e-4, -4e-4, -0.01, 0.01, 0.002, 0.001, 1e-3, -2e-3, 5e-4, 3e-4, 0.02, -0.015])
RVEC = np.array([0.01, -0.02, 0.005])


def points():
    return np.random.default_rng(3).uniform([100, 50], [1180, 670], (500, 2))


if __name__ == "__main__":
    import cv2
    pts = points()
    R = cv2.Rodrigues(RVEC)[0]
    np.savez_compressed(os.path.join(HERE, "golden_tilt.npz"), R=R,
                        plain=cv2.undistortPoints(pts.reshape(-1, 1, 2), K, DIST),
                        rect=cv2.undistortPoints(pts.reshape(-1, 1, 2), K, DIST, R=R, P=K),
                        f32=cv2.undistortPoints(pts.astype(np.float32).reshape(-1, 1, 2), K, DIST, R=R, P=K))
    print("wrote golden_tilt.npz, cv2", cv2.__version__)
