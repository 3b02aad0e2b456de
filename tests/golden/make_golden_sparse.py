"""Golden vectors for the sparse 3-D path + ingest, generated from the cv2 4.13.0 binary of this
container.  Run here (needs cv2):  python tests/golden/make_golden_sparse.py  -> tests/golden/golden_sparse.npz"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from stereo_vision_b200 import synth  # noqa: E402


def rodrigues(v):
    v = np.asarray(v, np.float64)
    th = np.linalg.norm(v)
    k = v / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)


def rig():
    K1, d1, K2, d2, R, T = synth.synthetic_rig(1920, 1080)
    return K1, np.asarray(d1, np.float64), K2, np.asarray(d2, np.float64), R, np.asarray(T, np.float64).reshape(3, 1)


def cases():
    """Seeded inputs shared by the generator and the tests."""
    rng = np.random.default_rng(20251104)
    K1, d1, K2, d2, R, T = rig()
    out = {"K1": K1, "d1": d1, "K2": K2, "d2": d2, "R": R, "T": T}
    d8 = np.concatenate([d1, [0.01, -0.002, 0.0003]])
    d12 = np.concatenate([d8, [1e-4, -2e-4, 3e-4, 1e-4]])
    Rr = rodrigues([0.01, -0.02, 0.005])
    P34 = np.array([[1300., 0, 960, 5], [0, 1300, 540, 1], [0, 0, 1, 0]])
    out.update(d8=d8, d12=d12, Rr=Rr, P34=P34)
    out["pts32"] = (rng.random((257, 1, 2)) * [1920, 1080]).astype(np.float32)
    out["pts64"] = (rng.random((1, 64, 2)) * [1920, 1080]).astype(np.float64)
    X = rng.random((300, 3)) * [2000, 1000, 3000] + [-1000, -500, 2000]
    P1 = K1 @ np.hstack([np.eye(3), np.zeros((3, 1))])
    P2 = K2 @ np.hstack([R, T])

    def proj(P):
        h = P @ np.vstack([X.T, np.ones(len(X))])
        return h[:2] / h[2]
    out.update(P1=P1, P2=P2, X=X, a64=proj(P1) + rng.normal(0, 0.3, (2, 300)), b64=proj(P2) + rng.normal(0, 0.3, (2, 300)))
    out["a32"], out["b32"] = out["a64"].astype(np.float32), out["b64"].astype(np.float32)
    out["bgr"] = rng.integers(0, 256, (97, 131, 3), dtype=np.uint8)
    out["bgr_small"] = synth.stereo_pair(120, 200, 32, seed=7, channels=3)[0]
    out["lp"] = (rng.random((40, 2)) * [1400, 800] + [260, 140]).astype(np.float32)
    out["rp"] = out["lp"] - np.array([60.0, 0.5], np.float32)
    # small rig for the loader / rectify glue (checkerboard_distance.py:49-103)
    k1, e1, k2, e2, r, t = synth.synthetic_rig(320, 180)
    out["rig"] = {"left_matrix": k1, "left_dist": np.asarray(e1, np.float64), "right_matrix": k2,
                  "right_dist": np.asarray(e2, np.float64), "R": r, "T": np.asarray(t, np.float64).reshape(3, 1)}
    out["rig_left"], out["rig_right"] = synth.stereo_pair(180, 320, 32, seed=11, channels=3)
    out["rig_lp"] = (rng.random((54, 2)) * [200, 120] + [60, 30]).astype(np.float32)
    out["rig_rp"] = out["rig_lp"] - np.array([9.0, 0.2], np.float32)
    return out


def main():
    import cv2
    c = cases()
    g = {}
    for name, d in (("d5", c["d1"]), ("d8", c["d8"]), ("d12", c["d12"]), ("d0", np.zeros(5))):
        g[f"und32_{name}_PK"] = cv2.undistortPoints(c["pts32"], c["K1"], d, P=c["K1"])
        g[f"und64_{name}_none"] = cv2.undistortPoints(c["pts64"], c["K1"], d)
        g[f"und32_{name}_RP"] = cv2.undistortPoints(c["pts32"], c["K1"], d, R=c["Rr"], P=c["P34"])
    g["tri64"] = cv2.triangulatePoints(c["P1"], c["P2"], c["a64"], c["b64"])
    g["tri32"] = cv2.triangulatePoints(c["P1"], c["P2"], c["a32"], c["b32"])
    g["gray"] = cv2.cvtColor(c["bgr"], cv2.COLOR_BGR2GRAY)
    g["gray_small"] = cv2.cvtColor(c["bgr_small"], cv2.COLOR_BGR2GRAY)
    g["undist_img"] = cv2.undistort(c["bgr_small"], c["K1"] * [[0.1, 1, 0.1], [1, 0.1, 0.1], [1, 1, 1]], c["d1"])
    Ks = c["K1"] * [[0.1, 1, 0.1], [1, 0.1, 0.1], [1, 1, 1]]
    g["fix_map1"], g["fix_map2"] = cv2.initUndistortRectifyMap(Ks, c["d1"], None, Ks, (200, 120), cv2.CV_16SC2)
    # the reference's calculate_3d_position (ball_drop.py:225-259), point by point, with cv2
    P1 = c["K1"] @ np.hstack((np.eye(3), np.zeros((3, 1))))
    P2 = c["K2"] @ np.hstack((c["R"], c["T"]))
    pos = []
    for lp, rp in zip(c["lp"], c["rp"]):
        lu = cv2.undistortPoints(np.array([[lp]], np.float32), c["K1"], c["d1"], P=c["K1"])[0, 0]
        ru = cv2.undistortPoints(np.array([[rp]], np.float32), c["K2"], c["d2"], P=c["K2"])[0, 0]
        h = cv2.triangulatePoints(P1, P2, lu, ru)
        pos.append((h[:3] / h[3]).reshape(3))
    g["calc3d"] = np.array(pos)
    # loader / rectify glue: stereoRectify(alpha=0) -> maps -> remap; corners -> undistortPoints(R1,P1) -> triangulate
    rg = c["rig"]
    R1, R2, Pr1, Pr2, Q, _, _ = cv2.stereoRectify(rg["left_matrix"], rg["left_dist"], rg["right_matrix"], rg["right_dist"],
                                                 (320, 180), rg["R"], rg["T"], alpha=0.0)
    g.update(rig_R1=R1, rig_R2=R2, rig_P1=Pr1, rig_P2=Pr2, rig_Q=Q)
    ml = cv2.initUndistortRectifyMap(rg["left_matrix"], rg["left_dist"], R1, Pr1, (320, 180), cv2.CV_32FC1)
    mr = cv2.initUndistortRectifyMap(rg["right_matrix"], rg["right_dist"], R2, Pr2, (320, 180), cv2.CV_32FC1)
    g["rig_rect_left"] = cv2.remap(c["rig_left"], ml[0], ml[1], cv2.INTER_LINEAR)
    g["rig_rect_right"] = cv2.remap(c["rig_right"], mr[0], mr[1], cv2.INTER_LINEAR)
    lu = cv2.undistortPoints(c["rig_lp"].reshape(-1, 1, 2), rg["left_matrix"], rg["left_dist"], R=R1, P=Pr1)
    ru = cv2.undistortPoints(c["rig_rp"].reshape(-1, 1, 2), rg["right_matrix"], rg["right_dist"], R=R2, P=Pr2)
    pos = []
    for j in range(lu.shape[0]):
        h = cv2.triangulatePoints(Pr1, Pr2, lu[j, 0].reshape(2, 1), ru[j, 0].reshape(2, 1))
        pos.append((h[:3] / h[3]).reshape(3))
    g["rig_pts3d"] = np.array(pos)
    np.savez_compressed(os.path.join(HERE, "golden_sparse.npz"), **g)
    print("wrote", len(g), "arrays, cv2", cv2.__version__)


if __name__ == "__main__":
    main()
