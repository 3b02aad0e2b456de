"""Seeded case lists shared by make_golden.py (cv2 side, run once in the build container) and the
tests (oracle / CUDA side).  Inputs are regenerated from seeds; only cv2 OUTPUTS are stored."""
from __future__ import annotations

import numpy as np

from stereo_vision_b200 import synth


def sgbm_small_cases(n: int = 96, seed: int = 20261104, modes=(0, 1), hrange=(30, 44)):
    """Random sweep over the SURVEY.md §8c matrix; all in-contract (maxC + P2 <= 32767)."""
    rng = np.random.default_rng(seed)
    out = []
    it = 0
    while len(out) < n:
        it += 1
        mode = int(modes[int(rng.integers(0, len(modes)))]); minD = int(rng.choice([-4, 0, 4])); D = int(rng.choice([16, 24, 32, 48, 64]))
        bs = int(rng.choice([1, 3, 5, 7, 9, 11])); cap = int(rng.choice([0, 31, 63])); uq = int(rng.choice([0, 5, 10, 15]))
        d12 = int(rng.choice([0, 1, 2, 3])); ch = int(rng.choice([1, 3]))
        kind = str(rng.choice(["smooth", "noise", "binary"]))
        H = int(rng.integers(*hrange)); W = int(rng.integers(D + abs(minD) + 12, D + abs(minD) + 90))
        P1 = int(rng.choice([0, 8 * ch * bs * bs])); P2 = int(rng.choice([0, 32 * ch * bs * bs]))
        sw = int(rng.choice([0, 20, 100])); sr = int(rng.choice([1, 2]))
        kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=d12,
                  preFilterCap=cap, uniquenessRatio=uq, speckleWindowSize=sw, speckleRange=sr, mode=mode)
        ft = max(cap, 15) | 1
        r = (bs if bs > 0 else 5) // 2
        p1e = P1 if P1 > 0 else 2
        p2e = max(P2 if P2 > 0 else 5, p1e + 1)
        if (2 * r + 1) ** 2 * ch * (2 * ft + 63) + p2e > 32767:
            continue
        out.append(dict(kw=kw, H=H, W=W, ch=ch, kind=kind, seed=it))
    return out


def sgbm_hh4_cases(n: int = 40):
    """Same matrix for MODE_HH4 (mode 3)."""
    return sgbm_small_cases(n, seed=20261105, modes=(3,))


def sgbm_3way_cases(n: int = 48):
    """Same matrix for MODE_SGBM_3WAY (mode 2): taller images, so that the four row stripes and their overlaps matter."""
    return sgbm_small_cases(n, seed=20261106, modes=(2,), hrange=(40, 90))


def sgbm_inputs(case):
    return synth.stereo_pair(case["H"], case["W"], case["kw"]["numDisparities"], seed=case["seed"],
                             kind=case["kind"], channels=case["ch"])


# BASELINE.json configs (full size): name -> (H, W, D, bs, mode, speckle)
FULL = {
    "c1": dict(H=480, W=640, D=64, bs=5, mode=0, speckle=False),
    "c2": dict(H=720, W=1280, D=128, bs=5, mode=0, speckle=True),
    "c3": dict(H=1080, W=1920, D=256, bs=5, mode=1, speckle=True),
    "c3_sgbm": dict(H=1080, W=1920, D=256, bs=5, mode=0, speckle=True),
    "c5": dict(H=2160, W=3840, D=256, bs=7, mode=0, speckle=True),
}


def full_inputs(name, seed=1000):
    c = FULL[name]
    L, R = synth.stereo_pair(c["H"], c["W"], c["D"], seed=seed, kind="smooth")
    kw = synth.sgbm_kwargs(c["D"], c["bs"], c["mode"], speckle=c["speckle"])
    return L, R, kw


def remap_cases():
    return [dict(H=97, W=131, C=1, seed=1), dict(H=97, W=131, C=3, seed=2), dict(H=240, W=320, C=3, seed=3),
            dict(H=64, W=64, C=1, seed=4)]


def remap_inputs(c):
    rng = np.random.default_rng(c["seed"])
    H, W, C = c["H"], c["W"], c["C"]
    src = rng.integers(0, 256, (H, W, C), dtype=np.uint8)
    if C == 1:
        src = np.ascontiguousarray(src[:, :, 0])
    mx = (np.arange(W)[None, :] + rng.uniform(-40, 40, (H, W))).astype(np.float32)
    my = (np.arange(H)[:, None] + rng.uniform(-40, 40, (H, W))).astype(np.float32)
    mx[0, :8] = [np.nan, 1e9, -1e9, 3.015625, 5.046875, -0.015625, W - 1, W - 0.5]
    my[1, :4] = [np.nan, 1e9, -1e9, 2.5 + 1 / 64]
    mx[2, :] = np.arange(W) + 1 / 64 * (np.arange(W) % 64)   # 1/64-step tie coordinates
    my[2, :] = 2 + 1 / 64
    return src, mx, my


def post_inputs(seed=9):
    """Noisy/blocky int16 disparity-like maps for median3 / filterSpeckles."""
    rng = np.random.default_rng(seed)
    a = (rng.integers(0, 20, (60, 80)) * 16).astype(np.int16)
    a[rng.random((60, 80)) < 0.2] = -16
    b = (np.kron(rng.integers(0, 12, (6, 8)), np.ones((10, 10))).astype(np.int16) * 16
         + rng.integers(-8, 9, (60, 80)).astype(np.int16))
    return a, b


SPECKLE_PARAMS = [(20, 16), (100, 32), (16, 0), (5, 200)]


# ---- round 2 (make_golden_r2.py) ----
REF_V5_SIZE = (1080, 1920)   # (W, H): the calibrate_v5 cameras' principal point is ~ (550, 500); portrait 1080p


def ref_v5_frame(seed=41):
    """A random BGR frame of the fixture's size (smoothed so that bilinear weights matter)."""
    W, H = REF_V5_SIZE
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray(np.stack([synth._gauss_blur_u8(rng.integers(0, 256, (H, W), dtype=np.uint8))
                                          for _ in range(3)], -1))


def chain_c2_inputs(seed=2000):
    """Raw (unrectified) 1280x720 BGR pair for the c2 pipeline chain."""
    return synth.stereo_pair(720, 1280, 128, seed=seed, channels=3)
