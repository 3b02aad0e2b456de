"""Workload for compute-sanitizer (tools/gpu_sanitize.sh): c1 (640x480, D=64, MODE_SGBM), a 3-band x 4-frame MODE_HH
batch at Dp=256 (cross-band hand-off, ticket scheduler, margin prefetches), MODE_HH4, MODE_SGBM_3WAY stripes, 3-channel
input, then remap / reproject / box medians / speckle.  Every result is checked against the oracle."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import oracle
import stereo_vision_b200 as sv
from stereo_vision_b200 import synth


def main():
    L, R = synth.stereo_pair(480, 640, 64, seed=1000)
    kw = synth.sgbm_kwargs(64, 5, 0, speckle=True)
    d = sv.StereoSGBM_create(**kw).compute(L, R)
    assert np.array_equal(d, oracle.sgbm_compute(L, R, **kw)), "c1"
    print("c1 ok", flush=True)
    H, W, D, B = 45, 400, 256, 4
    for mode in (1, 3, 2):
        if mode == 2:
            H = 64
        kw = synth.sgbm_kwargs(D, 5, mode)
        pairs = [synth.stereo_pair(H, W, D, seed=10 + i) for i in range(B)]
        Lb = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
        Rb = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
        out = sv.StereoSGBM_create(**kw).compute_batch(Lb, Rb).cpu().numpy()
        for i, (l, r) in enumerate(pairs):
            assert np.array_equal(out[i], oracle.sgbm_compute(l, r, **kw)), (mode, i)
        print("mode", mode, "batch ok", flush=True)
    L3, R3 = synth.stereo_pair(40, 200, 96, seed=5, channels=3)
    kw = synth.sgbm_kwargs(96, 5, 1, channels=3)
    assert np.array_equal(sv.StereoSGBM_create(**kw).compute(L3, R3), oracle.sgbm_compute(L3, R3, **kw)), "3ch"
    print("3ch ok", flush=True)
    W, H = 320, 240
    K1, d1, K2, d2, Rm, T = synth.synthetic_rig(W, H)
    R1, R2, P1, P2, Q = synth.simple_rectify(K1, K2, Rm, T, W, H)
    mx, my = sv.initUndistortRectifyMap(K1, d1, R1, P1, (W, H), sv.CV_32FC1)
    src = np.random.default_rng(0).integers(0, 256, (H, W, 3), dtype=np.uint8)
    ox, oy = oracle.init_undistort_rectify_map(K1, d1, R1, P1, (W, H))
    assert np.array_equal(sv.remap(src, mx, my, sv.INTER_LINEAR), oracle.remap(src, ox, oy)), "remap"
    disp = np.random.default_rng(1).integers(-16, 64 * 16, (H, W)).astype(np.int16)
    assert np.array_equal(sv.reprojectImageTo3D(disp, Q), oracle.reproject_to_3d(disp, Q), equal_nan=True), "reproject"
    boxes = synth.boxes_for_frame(H, W, 32, 4, seed=0)
    assert np.array_equal(sv.box_median_depth(disp, Q, boxes), oracle.box_median_depth(disp, Q, boxes, -16), equal_nan=True)
    a = disp.copy()
    sv.filterSpeckles(a, -16, 50, 32)
    assert np.array_equal(a, oracle.filter_speckles(disp, -16, 50, 32)), "speckle"
    torch.cuda.synchronize()
    print("all ok", flush=True)


if __name__ == "__main__":
    main()
