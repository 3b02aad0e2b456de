"""Ad-hoc GPU bring-up: stage-by-stage diff against the oracle + rough timings."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
import numpy as np, torch
import oracle, cases
import stereo_vision_b200 as sv
from stereo_vision_b200 import synth

def stage(H, W, D, bs, ch, minD, kind, mode):
    L, R = synth.stereo_pair(H, W, D, seed=3, kind=kind, channels=ch)
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8*ch*bs*bs, P2=32*ch*bs*bs, preFilterCap=63, uniquenessRatio=10, mode=mode)
    raw_o, C_o, S_o = oracle.sgbm_stages(L, R, **kw)
    try:
        C_g, S_g, raw_g = sv.StereoSGBM_create(**kw).debug_stages(L, R)
    except Exception as e:
        print("EXC", H, W, D, bs, ch, minD, kind, mode, repr(e)); return
    r = oracle.resolve(**kw)
    S4 = oracle.aggregate(C_o, r["P1"], r["P2"], 4)
    cd = (C_g[:, :, :D] != C_o); sd = (S_g[:, :, :D] != S4); rd = raw_g != raw_o
    print(f"H{H} W{W} D{D} bs{bs} ch{ch} minD{minD} {kind} mode{mode}: C diff {cd.sum()} S4 diff {sd.sum()} raw diff {rd.sum()}")
    if cd.any():
        idx = np.argwhere(cd)[:5]; print("  C first diffs", idx.tolist(), [(int(C_g[tuple(i)]), int(C_o[tuple(i)])) for i in idx])
    elif sd.any():
        idx = np.argwhere(sd)[:5]; print("  S first diffs", idx.tolist(), [(int(S_g[tuple(i)]), int(S4[tuple(i)])) for i in idx])
    elif rd.any():
        idx = np.argwhere(rd)[:5]; print("  raw first diffs", idx.tolist(), [(int(raw_g[tuple(i)]), int(raw_o[tuple(i)])) for i in idx])
    full_o = oracle.sgbm_compute(L, R, **kw, speckleWindowSize=50, speckleRange=2)
    full_g = sv.StereoSGBM_create(**kw, speckleWindowSize=50, speckleRange=2).compute(L, R)
    print("   full diff", (full_o != full_g).sum())

print(torch.cuda.get_device_name(0))
for mode in (0, 1):
    stage(40, 150, 64, 5, 1, 0, "smooth", mode)
    stage(37, 120, 24, 3, 3, -4, "noise", mode)
    stage(33, 100, 16, 11, 1, 4, "binary", mode)
    stage(36, 340, 256, 7, 1, 0, "smooth", mode)
    stage(70, 200, 128, 1, 1, 0, "smooth", mode)

def timeit(name, H, W, D, bs, mode, B=1, n=5):
    L, R = synth.stereo_pair(H, W, D, seed=1000)
    kw = synth.sgbm_kwargs(D, bs, mode)
    m = sv.StereoSGBM_create(**kw)
    Lb = torch.from_numpy(L).cuda()[None].repeat(B, 1, 1).contiguous(); Rb = torch.from_numpy(R).cuda()[None].repeat(B, 1, 1).contiguous()
    out = m.compute_batch(Lb, Rb); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): out = m.compute_batch(Lb, Rb)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name}: B={B} {ms:.3f} ms/batch  {ms/B:.3f} ms/frame  {B*1000/ms:.1f} fps  {W*H*D*B/ms/1e6:.1f} Gpix.disp/s", flush=True)
    return out

for B in (1, 4, 8):
    timeit("c1 640x480 D64 SGBM", 480, 640, 64, 5, 0, B)
for B in (1, 2, 4, 8):
    timeit("c3 1080p D256 HH", 1080, 1920, 256, 5, 1, B)
    timeit("c3 1080p D256 SGBM", 1080, 1920, 256, 5, 0, B)
import json, hashlib
gf = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests/golden/golden_full.json")))
for name in ("c1", "c2", "c3", "c3_sgbm"):
    L, R, kw = cases.full_inputs(name)
    got = sv.StereoSGBM_create(**kw).compute(L, R)
    print(name, "golden match:", hashlib.sha256(got.tobytes()).hexdigest() == gf["full"][name]["sha256"], flush=True)
