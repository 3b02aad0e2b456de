"""Per-stage device times of StereoSGBM.compute at the BASELINE.json configs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stereo_vision_b200 as sv
from stereo_vision_b200 import synth

def run(name, H, W, D, bs, mode, B, n=3):
    kw = synth.sgbm_kwargs(D, bs, mode)
    m = sv.StereoSGBM_create(**kw)
    Ls, Rs = zip(*[synth.stereo_pair(H, W, D, seed=1000 + i) for i in range(min(B, 2))])
    Lb = torch.from_numpy(np.stack([Ls[i % len(Ls)] for i in range(B)])).cuda()
    Rb = torch.from_numpy(np.stack([Rs[i % len(Rs)] for i in range(B)])).cuda()
    sv.set_stage_timing(True)
    m.compute_batch(Lb, Rb); torch.cuda.synchronize()
    acc = None
    for _ in range(n):
        m.compute_batch(Lb, Rb)
        t = sv.last_stage_ms()
        acc = t if acc is None else {k: acc[k] + t[k] for k in t}
    tot = sum(acc.values()) / n
    print(f"{name} B={B}: total {tot:.3f} ms ({tot/B:.3f}/frame) " + " ".join(f"{k}={v/n:.3f}" for k, v in acc.items()), flush=True)

if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    if B:
        run("c3 HH", 1080, 1920, 256, 5, 1, B, n=2)
    else:
        for B in (1, 8):
            run("c1 SGBM", 480, 640, 64, 5, 0, B)
            run("c2 SGBM", 720, 1280, 128, 5, 0, B)
            run("c3 HH", 1080, 1920, 256, 5, 1, B)
            run("c3 SGBM", 1080, 1920, 256, 5, 0, B)
        run("c5 4K SGBM", 2160, 3840, 256, 7, 0, 1)
