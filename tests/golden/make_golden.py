"""Generate golden vectors from the cv2 4.13.0 binary (the reference's un-vendored dependency).

Run ONCE in the build container:   python tests/golden/make_golden.py
Writes tests/golden/golden_small.npz (cv2 outputs for the small seeded cases) and
tests/golden/golden_full.json (SHA-256 of cv2 outputs at the BASELINE.json full sizes).
The GPU box never needs cv2: tests regenerate the inputs from the seeds in cases.py.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import cv2  # noqa: E402

import cases  # noqa: E402
from stereo_vision_b200 import synth  # noqa: E402


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    out = {}
    meta = {"cv2": cv2.__version__, "numpy": np.__version__}
    # --- SGBM small sweep
    for i, c in enumerate(cases.sgbm_small_cases()):
        L, R = cases.sgbm_inputs(c)
        out[f"sgbm_{i}"] = cv2.StereoSGBM_create(**c["kw"]).compute(L, R)
        out[f"sgbm_in_{i}"] = np.frombuffer(bytes.fromhex(sha(L)[:16] + sha(R)[:16]), np.uint8)
    # --- parameter-resolution equalities are tested through the outputs above + test_params
    # --- remap
    for i, c in enumerate(cases.remap_cases()):
        src, mx, my = cases.remap_inputs(c)
        out[f"remap_{i}"] = cv2.remap(src, mx, my, cv2.INTER_LINEAR)
    # --- maps on synthetic rigs: store the x32-quantised maps of a 160-row slab + full hash
    for (W, H) in [(640, 480), (1280, 720), (1920, 1080)]:
        K1, d1, K2, d2, R, T = synth.synthetic_rig(W, H)
        R1, R2, P1, P2, Q, _, _ = cv2.stereoRectify(K1, d1, K2, d2, (W, H), R, T, alpha=0)
        out[f"rect_{W}"] = np.concatenate([R1.ravel(), R2.ravel(), P1.ravel(), P2.ravel(), Q.ravel()])
        for side, (K, d, Rr, P) in enumerate([(K1, d1, R1, P1), (K2, d2, R2, P2)]):
            mx, my = cv2.initUndistortRectifyMap(K, d, Rr, P, (W, H), cv2.CV_32FC1)
            qx = np.rint(mx * np.float32(32)).astype(np.int32)
            qy = np.rint(my * np.float32(32)).astype(np.int32)
            meta[f"mapq_{W}_{side}"] = [sha(qx), sha(qy)]
            if W == 640:
                out[f"mapx_{W}_{side}"] = mx[::8, ::8].copy()
                out[f"mapy_{W}_{side}"] = my[::8, ::8].copy()
    # --- reproject (640x480 rig Q), int16 and f32, with/without handleMissingValues; subsampled
    W, H = 640, 480
    rng = np.random.default_rng(5)
    Q = out["rect_640"][-16:].reshape(4, 4)
    disp = rng.integers(-16, 64 * 16, (H, W)).astype(np.int16)
    for hm in (0, 1):
        out[f"reproj_s16_{hm}"] = cv2.reprojectImageTo3D(disp, Q, handleMissingValues=bool(hm))[::4, ::4].copy()
        out[f"reproj_f32_{hm}"] = cv2.reprojectImageTo3D(disp.astype(np.float32) / 16, Q,
                                                         handleMissingValues=bool(hm))[::4, ::4].copy()
    # --- median / speckle
    for j, arr in enumerate(cases.post_inputs()):
        out[f"median_{j}"] = cv2.medianBlur(arr, 3)
        for (sz, df) in cases.SPECKLE_PARAMS:
            r = arr.copy()
            cv2.filterSpeckles(r, -16, sz, df)
            out[f"speckle_{j}_{sz}_{df}"] = r
    np.savez_compressed(os.path.join(HERE, "golden_small.npz"), **out)
    # --- full-size SGBM hashes (cv2 takes 0.07-9 s each)
    full = {}
    for name in cases.FULL:
        L, R, kw = cases.full_inputs(name)
        t0 = time.time()
        d = cv2.StereoSGBM_create(**kw).compute(L, R)
        full[name] = dict(sha256=sha(d), in_sha256=[sha(L), sha(R)], cv2_seconds=round(time.time() - t0, 3),
                          valid_frac=float((d >= 0).mean()), kw=kw)
        print(name, full[name]["cv2_seconds"], full[name]["valid_frac"], flush=True)
    meta["full"] = full
    with open(os.path.join(HERE, "golden_full.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
