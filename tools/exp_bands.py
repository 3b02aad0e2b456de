"""Experiment: sweep step time without cross-band coupling (single-band frames) and at larger batches."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stereo_vision_b200 as sv
from stereo_vision_b200 import synth
def run(name, H, W, D, bs, mode, B, n=2):
    kw = synth.sgbm_kwargs(D, bs, mode)
    m = sv.StereoSGBM_create(**kw)
    L, R = synth.stereo_pair(H, W, D, seed=1000)
    Lb = torch.from_numpy(L).cuda()[None].repeat(B, 1, 1).contiguous(); Rb = torch.from_numpy(R).cuda()[None].repeat(B, 1, 1).contiguous()
    sv.set_stage_timing(True)
    m.compute_batch(Lb, Rb); torch.cuda.synchronize()
    acc = None
    for _ in range(n):
        m.compute_batch(Lb, Rb)
        t = sv.last_stage_ms()
        acc = t if acc is None else {k: acc[k] + t[k] for k in t}
    tot = sum(acc.values()) / n
    print(f"{name} H={H} B={B}: total {tot:.3f} ms ({tot/B*1080/H:.3f}/1080-row frame) " + " ".join(f"{k}={v/n:.3f}" for k, v in acc.items()), flush=True)
    sv.release_workspaces()
for B in (148, 592, 1184):
    run("single band", 15, 1920, 256, 5, 1, B)
for B in (16, 32):
    run("c3", 1080, 1920, 256, 5, 1, B)
