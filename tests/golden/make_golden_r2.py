"""Round-2 golden vectors from the cv2 4.13.0 binary.  Run ONCE in the build container:

    python tests/golden/make_golden_r2.py

Writes tests/golden/golden_r2.json + golden_r2.npz:
  * ref_v5:  maps / remap on the reference-held calibration fixture
             /root/reference/calibration_tests/calibrate_v5/{calibration_results,stereo_calibration_results}
             (negative focal length in rect_P1, 5-coefficient distortion).  The ~40 numbers of the fixture are stored in
             the json so that the GPU box (which has no /root/reference) can rebuild the inputs.
  * chain_c2: raw 1280x720 BGR pair -> remap x2 -> cvtColor x2 -> StereoSGBM.compute -> reprojectImageTo3D -> box
             medians, every link by cv2 (hashes; XYZ subsampled; medians).
  * c3_3ch, c3_minD-16: full-size cv2 hashes for 3-channel input and minDisparity = -16.
  * c3_batch: cv2 hashes of the 32 distinct c3 frames bench.py runs per rank (seeds 1000..1031).
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import cv2  # noqa: E402

import cases  # noqa: E402
from stereo_vision_b200 import synth  # noqa: E402

REF = "/root/reference/calibration_tests/calibrate_v5"


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def qmap(m):
    return np.rint(m * np.float32(32)).astype(np.int32)


def ref_v5(meta, out):
    fx = {}
    for cam in (1, 2):
        fx[f"K{cam}"] = np.loadtxt(f"{REF}/calibration_results/cam{cam}_matrix.txt")
        fx[f"d{cam}"] = np.loadtxt(f"{REF}/calibration_results/cam{cam}_distortion.txt").ravel()
        fx[f"R{cam}"] = np.loadtxt(f"{REF}/stereo_calibration_results/rect_R{cam}.txt")
        fx[f"P{cam}"] = np.loadtxt(f"{REF}/stereo_calibration_results/rect_P{cam}.txt")
    W, H = cases.REF_V5_SIZE
    m = {"fixture": {k: v.tolist() for k, v in fx.items()}, "size": [W, H], "source": REF}
    src = cases.ref_v5_frame()
    for cam in (1, 2):
        mx, my = cv2.initUndistortRectifyMap(fx[f"K{cam}"], fx[f"d{cam}"], fx[f"R{cam}"], fx[f"P{cam}"], (W, H), cv2.CV_32FC1)
        m[f"mapq_{cam}"] = [sha(qmap(mx)), sha(qmap(my))]
        m[f"map_inside_frac_{cam}"] = float(((mx >= 0) & (mx < W) & (my >= 0) & (my < H)).mean())
        r = cv2.remap(src, mx, my, cv2.INTER_LINEAR)
        m[f"remap_{cam}"] = sha(r)
        m[f"remap_nonzero_frac_{cam}"] = float((r != 0).mean())
        out[f"ref_v5_mapx_{cam}"] = mx[::16, ::16].copy()
        out[f"ref_v5_mapy_{cam}"] = my[::16, ::16].copy()
        # the same distortion / rotation with a sane projection (P = K): the maps stay inside the image, so the remap
        # link is exercised on real pixels too
        mx, my = cv2.initUndistortRectifyMap(fx[f"K{cam}"], fx[f"d{cam}"], fx[f"R{cam}"], fx[f"K{cam}"], (W, H), cv2.CV_32FC1)
        m[f"mapq_K_{cam}"] = [sha(qmap(mx)), sha(qmap(my))]
        r = cv2.remap(src, mx, my, cv2.INTER_LINEAR)
        m[f"remap_K_{cam}"] = sha(r)
        m[f"remap_K_nonzero_frac_{cam}"] = float((r != 0).mean())
    meta["ref_v5"] = m


def chain_c2(meta, out):
    W, H, D = 1280, 720, 128
    L, R = cases.chain_c2_inputs()
    K1, d1, K2, d2, Rm, T = synth.synthetic_rig(W, H)
    R1, R2, P1, P2, Q, _, _ = cv2.stereoRectify(K1, d1, K2, d2, (W, H), Rm, T, alpha=0)
    m1 = cv2.initUndistortRectifyMap(K1, d1, R1, P1, (W, H), cv2.CV_32FC1)
    m2 = cv2.initUndistortRectifyMap(K2, d2, R2, P2, (W, H), cv2.CV_32FC1)
    Lr = cv2.remap(L, m1[0], m1[1], cv2.INTER_LINEAR)
    Rr = cv2.remap(R, m2[0], m2[1], cv2.INTER_LINEAR)
    Lg, Rg = cv2.cvtColor(Lr, cv2.COLOR_BGR2GRAY), cv2.cvtColor(Rr, cv2.COLOR_BGR2GRAY)
    kw = synth.sgbm_kwargs(D, 5, 0)
    disp = cv2.StereoSGBM_create(**kw).compute(Lg, Rg)
    xyz = cv2.reprojectImageTo3D(disp.astype(np.float32) / 16, Q)
    boxes = synth.boxes_for_frame(H, W, D, 8, seed=5)
    med = []
    for (x1, y1, x2, y2) in boxes:
        sl = (slice(int(np.floor(y1)), int(np.ceil(y2))), slice(int(np.floor(x1)), int(np.ceil(x2))))
        z = xyz[..., 2][sl]
        v = z[(disp[sl] != -16) & np.isfinite(z)]
        med.append(float(np.median(v)) if v.size else float("nan"))
    meta["chain_c2"] = dict(in_sha=[sha(L), sha(R)], rect=[sha(Lr), sha(Rr)], gray=[sha(Lg), sha(Rg)], disp=sha(disp),
                            valid_frac=float((disp >= 0).mean()), kw=kw, rectify=dict(R1=R1.tolist(), R2=R2.tolist(),
                            P1=P1.tolist(), P2=P2.tolist(), Q=Q.tolist()))
    out["chain_c2_xyz"] = xyz[::8, ::8].copy()
    out["chain_c2_med"] = np.array(med, np.float32)
    out["chain_c2_boxes"] = boxes


def c3_variants(meta):
    H, W, D = 1080, 1920, 256
    L3, R3 = synth.stereo_pair(H, W, D, seed=1000, channels=3)

    kw3 = synth.sgbm_kwargs(D, 5, 1, channels=3)
    d = cv2.StereoSGBM_create(**kw3).compute(L3, R3)
    meta["c3_3ch"] = dict(sha256=sha(d), in_sha256=[sha(L3), sha(R3)], kw=kw3, valid_frac=float((d >= 0).mean()))
    L, R = synth.stereo_pair(H, W, D, seed=1000)
    kwm = synth.sgbm_kwargs(D, 5, 1, minD=-16)
    d = cv2.StereoSGBM_create(**kwm).compute(L, R)
    meta["c3_minD-16"] = dict(sha256=sha(d), kw=kwm, valid_frac=float((d > -17 * 16).mean()))


def c3_batch(meta):
    H, W, D = 1080, 1920, 256
    kw = synth.sgbm_kwargs(D, 5, 1)
    cv2.setNumThreads(1)

    def one(seed):
        L, R = synth.stereo_pair(H, W, D, seed=seed)
        return sha(cv2.StereoSGBM_create(**kw).compute(L, R))
    with ThreadPoolExecutor(8) as ex:
        hs = list(ex.map(one, range(1000, 1032)))
    meta["c3_batch"] = dict(seeds=[1000, 1032], kw=kw, sha256=hs)


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    meta, out = {"cv2": cv2.__version__}, {}
    ref_v5(meta, out)
    print("ref_v5 done", flush=True)
    chain_c2(meta, out)
    print("chain done", flush=True)
    c3_variants(meta)
    print("c3 variants done", flush=True)
    c3_batch(meta)
    np.savez_compressed(os.path.join(HERE, "golden_r2.npz"), **out)
    with open(os.path.join(HERE, "golden_r2.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
