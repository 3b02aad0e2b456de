"""BASELINE.json configs[1] ("c2"): 1280x720 BGR raw pairs, numDisparities=128 — the full rectify -> SGBM ->
reprojectImageTo3D pipeline (+ BGR2GRAY ingest and per-box median depth) on one B200, per-stage device times from CUDA
events on the launching stream and, for the HBM-bound stages, achieved algorithmic GB/s against the
measured copy peak.  Also times the 1080p BGR remap / reproject / maps (the sizes §8a quotes for a1, a2, a9).

    python tools/pipeline_c2.py [B] [--json out.json]
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stereo_vision_b200 as sv
from stereo_vision_b200 import synth


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def stages(H, W, D, B, tag, out):
    peak = peak_gbs()
    K1, d1, K2, d2, Rm, T = synth.synthetic_rig(W, H)
    R1, R2, P1, P2, Q = synth.simple_rectify(K1, K2, Rm, T, W, H)
    Ls, Rs = zip(*[synth.stereo_pair(H, W, D, seed=1000 + i) for i in range(min(B, 2))])
    bgr = lambda g: np.repeat(g[..., None], 3, axis=2)
    Lb = torch.from_numpy(np.stack([bgr(Ls[i % len(Ls)]) for i in range(B)])).cuda()   # raw BGR frames [B,H,W,3]
    Rb = torch.from_numpy(np.stack([bgr(Rs[i % len(Rs)]) for i in range(B)])).cuda()
    res = {}

    def rec(name, ms, alg_bytes=None):
        res[name] = {"ms": ms}
        line = f"{tag} B={B} {name:28s} {ms:8.3f} ms"
        if alg_bytes:
            g = alg_bytes / ms / 1e6
            res[name].update(algorithmic_bytes=alg_bytes, achieved_gbs=g, frac_of_hbm_peak=g / peak)
            line += f"  {alg_bytes/1e6:9.1f} MB algorithmic -> {g:7.0f} GB/s = {100*g/peak:4.1f}% of {peak:.0f}"
        print(line, flush=True)

    maps = {}
    def mk_maps():
        maps["l"] = sv.initUndistortRectifyMap(K1, d1, R1, P1, (W, H), sv.CV_32FC1, device="cuda")
        maps["r"] = sv.initUndistortRectifyMap(K2, d2, R2, P2, (W, H), sv.CV_32FC1, device="cuda")
    rec("initUndistortRectifyMap x2", timeit(mk_maps), 2 * 8.0 * W * H)                       # 2 f32 maps per camera
    rect = {}
    def do_remap():
        rect["l"] = sv.remap(Lb, *maps["l"], sv.INTER_LINEAR)
        rect["r"] = sv.remap(Rb, *maps["r"], sv.INTER_LINEAR)
    rec("remap BGR x2", timeit(do_remap), 2.0 * (8 + 2 * 3) * W * H * B)                     # maps 8 B + read 3 + write 3 per px
    gray = {}
    def do_gray():
        gray["l"] = sv.cvtColor(rect["l"], sv.COLOR_BGR2GRAY)
        gray["r"] = sv.cvtColor(rect["r"], sv.COLOR_BGR2GRAY)
    rec("BGR2GRAY x2", timeit(do_gray), 2.0 * 4 * W * H * B)
    m = sv.StereoSGBM_create(**synth.sgbm_kwargs(D, 5, 0))
    disp = {}
    def do_sgbm():
        disp["d"] = m.compute_batch(gray["l"], gray["r"])
    rec("StereoSGBM.compute (MODE_SGBM)", timeit(do_sgbm, n=5), 4.0 * W * H * B)                # read L,R gray + write int16
    xyz = {}
    def do_reproj():
        xyz["p"] = sv.reprojectImageTo3D(disp["d"], Q)
    rec("reprojectImageTo3D (int16)", timeit(do_reproj), (2 + 12.0) * W * H * B)
    boxes = torch.from_numpy(np.concatenate([synth.boxes_for_frame(H, W, D, 8, seed=i) for i in range(B)])).cuda()
    bi = torch.from_numpy(np.repeat(np.arange(B, dtype=np.int32), 8)).cuda()
    rec("box_median_depth (8/frame)", timeit(lambda: sv.box_median_depth(disp["d"], Q, boxes, box_image=bi)))
    def whole():
        do_remap(); do_gray(); do_sgbm(); do_reproj(); sv.box_median_depth(disp["d"], Q, boxes, box_image=bi)
    ms = timeit(whole, n=5)
    rec("pipeline (maps resident)", ms)
    res["frames_per_s"] = B / ms * 1e3
    res["gpix_disp_per_s"] = B / ms * 1e3 * W * H * D / 1e9
    print(f"{tag} B={B} pipeline: {res['frames_per_s']:.1f} frame pairs/s, {res['gpix_disp_per_s']:.1f} Gpix*disp/s", flush=True)
    out[f"{tag}_B{B}"] = res


if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 8
    out = {}
    stages(720, 1280, 128, 1, "c2", out)
    stages(720, 1280, 128, B, "c2", out)
    stages(1080, 1920, 256, B, "1080p", out)
    if "--json" in sys.argv:
        json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
