#!/usr/bin/env python
"""bench.py — 1080p D=256 SGBM throughput (frames/s, Gpix*disp/s) on N B200s vs the cv2 CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--config c1|c2|c3|c4|c5] [--impl ours|reference]

Default workload (BASELINE.json configs[2]/[3], "c3"/"c4"): synthetic rectified 1920x1080 gray pairs, numDisparities=256,
blockSize=5, MODE_HH (8 paths), LR check, uniqueness, 3x3 median, speckle filter, then per-YOLO-box median depth
(8 synthetic boxes per frame).  One "step" = one batch of B DISTINCT frame pairs per GPU through that hot path.  Frame
pairs are independent, so ranks shard the batch with no data-path collective (weak scaling; NCCL is used only
for the barrier / max-over-ranks timing).  --config selects another BASELINE.json configuration as the main workload
(c2 = the full raw-BGR -> remap -> gray -> SGBM -> reproject -> box-median pipeline); the default run also reports
c1, c2, c5 and the single-pair latencies in `other_configs` / `latency_ms`.

Prints ONE JSON line (rank 0).  See DESIGN.md §measurement for how each field is obtained.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NBOX = 8
# BASELINE.json configs: geometry, matcher mode, default batch per GPU, whether the rectification front end is part of it
CONFIGS = {
    "c1": dict(H=480, W=640, D=64, bs=5, mode=0, speckle=False, batch=64, pipeline=False,
               desc="c1: 640x480 gray pairs, numDisparities=64, blockSize=5, MODE_SGBM 5-path, LR check, 8 box medians/frame"),
    "c2": dict(H=720, W=1280, D=128, bs=5, mode=0, speckle=True, batch=32, pipeline=True,
               desc="c2: 1280x720 raw BGR pairs -> remap x2 -> BGR2GRAY x2 -> SGBM (numDisparities=128, blockSize=5, MODE_SGBM, "
                    "LR check + speckle) -> reprojectImageTo3D -> 8 box medians/frame"),
    "c3": dict(H=1080, W=1920, D=256, bs=5, mode=1, speckle=True, batch=32, pipeline=False,
               desc="c3/c4: 1920x1080 pairs, numDisparities=256, blockSize=5, MODE_HH 8-path, LR check + speckle, 8 box medians/frame"),
    "c5": dict(H=2160, W=3840, D=256, bs=7, mode=0, speckle=True, batch=4, pipeline=False,
               desc="c5: 3840x2160 gray pairs, numDisparities=256, blockSize=7, MODE_SGBM 5-path, LR check + speckle, 8 box medians/frame"),
}
CONFIGS["c4"] = CONFIGS["c3"]   # c4 = c3's workload sharded over 8 GPUs (256 pairs)
H, W, D, BS, MODE = 1080, 1920, 256, 5, 1
WORKLOAD = CONFIGS["c3"]["desc"]


def _sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _profile_numbers():
    """Numbers taken from the committed ncu captures (profiles/r02_counters.json, written by tools/ncu_summary.py from
    the --set full captures named inside): per-launch DRAM traffic and hardware pipe counters of the sweeps."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_counters.json")) as f:
            return json.load(f)
    except Exception:
        try:
            with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
                t = json.load(f)["k_sweep_pass2"]
            return {"k_sweep_pass2": {"dram_bytes": t["dram_bytes"], "batch": t["batch"], "source": t["source"]}}
        except Exception:
            return {}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_batch(cfg: dict, B: int, rank: int):
    """B DISTINCT seeded frame pairs per rank (seeds 1000 + 64*rank + i: rank 0's first 32 c3 frames are the ones whose
    cv2 hashes tests/golden/golden_r2.json holds).  c2 gets raw BGR frames (3 channels)."""
    from stereo_vision_b200 import synth
    Hc, Wc, Dc = cfg["H"], cfg["W"], cfg["D"]
    ch = 3 if cfg["pipeline"] else 1
    pairs = [synth.stereo_pair(Hc, Wc, Dc, seed=1000 + rank * 64 + i, channels=ch) for i in range(B)]
    L = np.stack([p[0] for p in pairs])
    R = np.stack([p[1] for p in pairs])
    boxes = np.concatenate([synth.boxes_for_frame(Hc, Wc, Dc, NBOX, seed=rank * 64 + i) for i in range(B)])
    box_image = np.repeat(np.arange(B, dtype=np.int32), NBOX)
    K1, d1, K2, d2, Rm, T = synth.synthetic_rig(Wc, Hc)
    R1, R2, P1, P2, Q = synth.simple_rectify(K1, K2, Rm, T, Wc, Hc)
    rig = dict(K1=K1, d1=d1, K2=K2, d2=d2, R1=R1, R2=R2, P1=P1, P2=P2)
    return L, R, boxes, box_image, Q, rig


def cpu_reference_fps(cfg: dict, n_threads: int, frames_per_thread: int = 1, form: str = "pool"):
    """Times the reference's own CPU implementation of the path on the host cores:
      form "pool"   (row ii): a thread pool over frame pairs, cv2.setNumThreads(1) inside (cv2's SGBM core is
                              single-threaded) — the throughput form;
      form "single" (row i) : cv2.setNumThreads(n_threads), one compute() at a time — the drop-in call as the reference's
                              scripts would make it.
    Falls back to the C oracle port when cv2 is not importable.  Returns (frames/s, kind, cores, sample description)."""
    from concurrent.futures import ThreadPoolExecutor
    from stereo_vision_b200 import synth
    Hc, Wc, Dc = cfg["H"], cfg["W"], cfg["D"]
    kw = synth.sgbm_kwargs(Dc, cfg["bs"], cfg["mode"], speckle=cfg["speckle"])
    ch = 3 if cfg["pipeline"] else 1
    L, R = synth.stereo_pair(Hc, Wc, Dc, seed=1000, channels=ch)
    K1, d1, K2, d2, Rm, T = synth.synthetic_rig(Wc, Hc)
    R1, R2, P1, P2, Q = synth.simple_rectify(K1, K2, Rm, T, Wc, Hc)
    boxes = synth.boxes_for_frame(Hc, Wc, Dc, NBOX, seed=0)
    try:
        import cv2
        kind = "reference"
        cv2.setNumThreads(1 if form == "pool" else n_threads)
        maps = None
        if cfg["pipeline"]:
            maps = (cv2.initUndistortRectifyMap(K1, d1, R1, P1, (Wc, Hc), cv2.CV_32FC1),
                    cv2.initUndistortRectifyMap(K2, d2, R2, P2, (Wc, Hc), cv2.CV_32FC1))

        def one(_):
            m = cv2.StereoSGBM_create(**kw)
            for _ in range(frames_per_thread):
                l, r = L, R
                if maps is not None:
                    l = cv2.cvtColor(cv2.remap(L, maps[0][0], maps[0][1], cv2.INTER_LINEAR), cv2.COLOR_BGR2GRAY)
                    r = cv2.cvtColor(cv2.remap(R, maps[1][0], maps[1][1], cv2.INTER_LINEAR), cv2.COLOR_BGR2GRAY)
                d = m.compute(l, r)
                z = cv2.reprojectImageTo3D(d.astype(np.float32) / 16, Q)[..., 2]
                for (x1, y1, x2, y2) in boxes:   # per-box median depth, as the GPU arm computes
                    sl = (slice(int(np.floor(y1)), int(np.ceil(y2))), slice(int(np.floor(x1)), int(np.ceil(x2))))
                    v = z[sl][(d[sl] != -16) & np.isfinite(z[sl])]
                    _ = np.median(v) if v.size else np.nan
        what = "rectify+gray+StereoSGBM.compute+reproject" if maps is not None else "StereoSGBM.compute"
        if form == "pool":
            desc = f"cv2 {cv2.__version__} {what}, pool of {n_threads} threads x {frames_per_thread} pairs ({cfg['desc'][:2]}), setNumThreads(1)"
            t0 = time.perf_counter()
            with ThreadPoolExecutor(n_threads) as ex:
                list(ex.map(one, range(n_threads)))
            dt = time.perf_counter() - t0
            nfr = n_threads * frames_per_thread
        else:
            desc = f"cv2 {cv2.__version__} {what}, one call at a time x {frames_per_thread} pairs ({cfg['desc'][:2]}), setNumThreads({n_threads})"
            t0 = time.perf_counter()
            one(0)
            dt = time.perf_counter() - t0
            nfr = frames_per_thread
    except ImportError:
        import oracle
        kind = "port"
        oracle.set_num_threads(n_threads)
        nfr = 1
        g = L if ch == 1 else oracle.bgr2gray(L)
        gr = R if ch == 1 else oracle.bgr2gray(R)
        desc = f"oracle/sv_oracle.c (OpenMP, {n_threads} threads), 1 pair ({cfg['desc'][:2]}), matcher only"
        t0 = time.perf_counter()
        oracle.sgbm_compute(g, gr, **kw)
        dt = time.perf_counter() - t0
    return nfr / dt, kind, n_threads, desc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    nthr = max(1, min(cores, 32))
    vals = []
    for i in range(args.warmup + args.steps):
        fps, kind, c, desc = cpu_reference_fps(cfg, nthr)
        if i >= args.warmup:
            vals.append(fps)
    v = float(np.mean(vals))
    line = {"impl": "reference", "metric": "frames_per_s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * nthr / v, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
            "gpix_disp_per_s": v * cfg["W"] * cfg["H"] * cfg["D"] / 1e9,
            "config": {"workload": cfg["desc"], "frames_per_step_per_gpu": nthr, "parallelism": f"host thread pool x{nthr}, no GPU",
                       "l2": "n/a (CPU arm)", "host_cores": cores},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": c, "kind": kind, "sample": desc},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def check_parity(name: str, wl) -> dict:
    """SHA-256 of the disparity maps the last timed end-to-end step left in host memory against cv2's hashes of the same
    seeded frames (tests/golden: golden_r2.json c3_batch = seeds 1000..1031; golden_full.json = seed 1000 of c1/c5)."""
    try:
        if name in ("c3", "c4"):
            with open(os.path.join(ROOT, "tests", "golden", "golden_r2.json")) as f:
                want = json.load(f)["c3_batch"]["sha256"]
            src = "tests/golden/golden_r2.json c3_batch (cv2 4.13.0, seeds 1000..)"
        elif name in ("c1", "c5"):
            with open(os.path.join(ROOT, "tests", "golden", "golden_full.json")) as f:
                want = [json.load(f)["full"][name]["sha256"]]
            src = f"tests/golden/golden_full.json full.{name} (cv2 4.13.0, seed 1000)"
        else:
            return {"ok": None, "frames": 0, "against": "none for the raw-BGR pipeline inputs: the chain is pinned by "
                    "tests/test_gpu_ops.py::test_chain_c2_pipeline_matches_cv2_link_by_link"}
        n = min(len(want), wl.B)
        bad = [i for i in range(n) if _sha(wl.disp_h[i].numpy()) != want[i]]
        return {"ok": not bad, "frames": n, "bad_frames": bad, "against": src}
    except Exception as e:   # a missing fixture must not kill the measurement, but it is not a pass either
        return {"ok": None, "frames": 0, "against": f"unavailable: {e}"}


def rooflines(name: str, wl, stage_ms: dict, W1: int, Dp: int, local: int) -> dict:
    """Roofline objects for the dominant kernel.  The sweeps are packed-int ALU-bound by design
    (north_star): `roofline` reports the dominant sweep against the packed s16x2 ALU peak measured live by
    svb_measure_packed_alu_peak; `hbm_roofline` is the same kernel against the measured HBM copy peak."""
    sv = wl.sv
    B, Hc, Dc = wl.B, wl.H, wl.D
    hh = wl.cfg["mode"] == 1
    dom = "aggregate_pass2_select" if hh else "aggregate_pass1"
    kern = ("k_sweep<NV=3,PASS2> (aggregation pass 2: 4 paths + S + selection)" if hh
            else "k_sweep<NV=3,!PASS2> (aggregation pass 1: 4 paths)")
    alg_bytes = 4.0 * W1 * Hc * Dp * B                      # 2 B of C + 2 B of S per (pixel, disparity)
    peak, peak_src = _peaks()
    ach = alg_bytes / (stage_ms[dom] / 1e3) / 1e9
    ops16_pass = 9.0 * W1 * Hc * Dc * 4 * B                # 9 16-bit ops per (px, d, path); 4 paths per pass
    alu = sv.packed_alu_peak(local)
    prof = _profile_numbers()
    pk = "k_sweep_pass2" if hh else "k_sweep_pass1"
    t = prof.get(pk, {})
    traffic = float(t["dram_bytes"]) * B / t["batch"] if (t and name in ("c3", "c4")) else None
    tsrc = ("%s (B=%d%s)" % (t.get("source"), t.get("batch"), "" if t.get("batch") == B else ", scaled x%g" % (B / t["batch"]))
            if traffic is not None else "no capture for this configuration")
    a_ops = ops16_pass / (stage_ms[dom] / 1e3)
    out = {
        "roofline": {"kernel": kern, "bound": "alu", "achieved": a_ops / 1e12, "peak": alu["ops16_per_s"] / 1e12,
                     "unit": "Tops16/s", "frac": a_ops / alu["ops16_per_s"], "traffic": traffic, "traffic_unit": "bytes/launch",
                     "traffic_source": tsrc, "peak_source": "measured live (svb_measure_packed_alu_peak: VIADDMNMX/VIMNMX.S16x2 chains on all SMs, x2 lanes)",
                     "algorithmic_ops16_per_launch": ops16_pass, "share_of_step": stage_ms[dom] / sum(stage_ms.values()),
                     "hw_pipe_alu_pct": t.get("pipe_alu_pct"), "hw_source": t.get("source")},
        "hbm_roofline": {"kernel": kern, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes},
    }
    if hh:
        agg_ms = stage_ms["aggregate_pass1"] + stage_ms[dom]
        t1 = prof.get("k_sweep_pass1", {})
        out["alu_roofline"] = {"stage": "aggregation (both sweeps)", "achieved_ops16_per_s": 2 * ops16_pass / (agg_ms / 1e3),
                               "peak_ops16_per_s": alu["ops16_per_s"], "frac": 2 * ops16_pass / (agg_ms / 1e3) / alu["ops16_per_s"],
                               "pass2_only_frac": a_ops / alu["ops16_per_s"],
                               "hw_pipe_alu_pct": {"pass1": t1.get("pipe_alu_pct"), "pass2": t.get("pipe_alu_pct"),
                                                   "source": t.get("source")},
                               "peak_source": "measured live (svb_measure_packed_alu_peak)"}
    # the cost kernel (second largest): C written once, against the HBM peak
    if "cost" in stage_ms and stage_ms["cost"] > 0:
        cb = 2.0 * W1 * Hc * Dp * B
        tc = prof.get("k_cost_fused", {})
        out["cost_roofline"] = {"kernel": "k_cost_fused", "bound": "hbm", "achieved": cb / (stage_ms["cost"] / 1e3) / 1e9, "peak": peak,
                                "unit": "GB/s", "frac": cb / (stage_ms["cost"] / 1e3) / 1e9 / peak, "algorithmic_bytes_per_launch": cb,
                                "hw_pipe_alu_pct": tc.get("pipe_alu_pct"), "hw_source": tc.get("source")}
    return out


def other_configs(local: int, timed) -> dict:
    """c1, c2 (full pipeline), c5 at their default batches (device-resident, 3 timed steps each) and the single-pair
    latency of every single-pair configuration: what `cv2.StereoSGBM_create(**kw).compute(L, R)` costs on the device."""
    import torch
    import stereo_vision_b200 as sv
    from stereo_vision_b200 import synth
    out, lat = {}, {}
    for name in ("c1", "c2", "c5"):
        cfg = CONFIGS[name]
        wl = Workload(name, cfg["batch"], 0, local)
        for _ in range(2):
            wl.step_resident()
        sv.set_stage_timing(True, local)
        acc = {}

        def collect():
            for k, v in sv.last_stage_ms(local).items():
                acc[k] = acc.get(k, 0.0) + v
        timed(wl.step_resident, 3, collect, sync=False)
        sv.set_stage_timing(False, local)
        ms = timed(wl.step_resident, 3, sync=False) / 3
        ms_e = timed(wl.step_e2e, 3, finish=wl.flush, sync=False) / 3
        torch.cuda.synchronize()
        fps = wl.B / ms * 1e3
        out[name] = {"workload": cfg["desc"], "frames_per_step": wl.B, "frames_per_s": fps, "ms_per_step": ms,
                     "gpix_disp_per_s": fps * cfg["W"] * cfg["H"] * cfg["D"] / 1e9, "e2e_frames_per_s": wl.B / ms_e * 1e3,
                     "stage_ms_per_step": {k: v / 3 for k, v in acc.items()}, "parity_checked": check_parity(name, wl)["ok"]}
        wl.close()
        del wl
        torch.cuda.empty_cache()
    for name in ("c1", "c2", "c3", "c5"):
        cfg = CONFIGS[name]
        L, R = synth.stereo_pair(cfg["H"], cfg["W"], cfg["D"], seed=1000)
        Ld, Rd = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
        m = sv.StereoSGBM_create(**synth.sgbm_kwargs(cfg["D"], cfg["bs"], cfg["mode"], speckle=cfg["speckle"]))
        for _ in range(3):
            m.compute(Ld, Rd)
        lat[name] = timed(lambda: m.compute(Ld, Rd), 5, sync=False) / 5
        sv.release_workspaces()
    return {"other_configs": out,
            "latency_ms": dict(lat, note="one rectified gray pair resident on the device through StereoSGBM.compute (B=1), CUDA events")}


class Workload:
    """One BASELINE.json configuration resident on one GPU: device-resident step, end-to-end step, output access."""

    def __init__(self, name: str, B: int, rank: int, local: int):
        import torch
        import stereo_vision_b200 as sv
        from stereo_vision_b200 import synth
        self.sv, self.torch = sv, torch
        self.name, self.cfg, self.B, self.local = name, CONFIGS[name], B, local
        cfg = self.cfg
        self.H, self.W, self.D = cfg["H"], cfg["W"], cfg["D"]
        dev = torch.device("cuda", local)
        L, R, boxes, box_image, self.Q, rig = make_batch(cfg, B, rank)
        self.kw = synth.sgbm_kwargs(self.D, cfg["bs"], cfg["mode"], speckle=cfg["speckle"])
        self.m = sv.StereoSGBM_create(**self.kw)
        self.Ld, self.Rd = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
        self.bx, self.bi = torch.from_numpy(boxes).to(dev), torch.from_numpy(box_image).to(dev)
        self.Lh, self.Rh = torch.from_numpy(L).pin_memory(), torch.from_numpy(R).pin_memory()   # pinned staging (e2e arm)
        self.disp_h = torch.empty((B, self.H, self.W), dtype=torch.int16).pin_memory()
        self.med_h = torch.empty(B * NBOX, dtype=torch.float32).pin_memory()
        self.disp_d = torch.empty((B, self.H, self.W), dtype=torch.int16, device=dev)
        self.pipeline = cfg["pipeline"]
        if self.pipeline:   # rectification maps stay resident per calibration (they are calibration-time data)
            size = (self.W, self.H)
            self.mapL = sv.initUndistortRectifyMap(rig["K1"], rig["d1"], rig["R1"], rig["P1"], size, sv.CV_32FC1, device=dev)
            self.mapR = sv.initUndistortRectifyMap(rig["K2"], rig["d2"], rig["R2"], rig["P2"], size, sv.CV_32FC1, device=dev)
            self.xyz_h = torch.empty((B, self.H, self.W, 3), dtype=torch.float32).pin_memory()
            self.Le, self.Re = torch.empty_like(self.Ld), torch.empty_like(self.Rd)
        else:
            self.pipe = sv.HostPipeline(self.m, device=local)
        # kernels per step: planes, cost, sweep x2, lrcheck, median3, [ccl x4], box_median (+ remap x2, gray x2, reproject)
        self.launches = 7 + (4 if cfg["speckle"] else 0) + (5 if self.pipeline else 0)
        ch = 3 if self.pipeline else 1
        self.h2d = int(2 * B * self.H * self.W * ch)
        self.d2h = int(B * self.H * self.W * 2 + B * NBOX * 4) + (int(B * self.H * self.W * 12) if self.pipeline else 0)

    def _front(self, Lraw, Rraw):
        sv = self.sv
        Lr, Rr = sv.remap(Lraw, *self.mapL, sv.INTER_LINEAR), sv.remap(Rraw, *self.mapR, sv.INTER_LINEAR)
        return sv.cvtColor(Lr, sv.COLOR_BGR2GRAY), sv.cvtColor(Rr, sv.COLOR_BGR2GRAY)

    def step_resident(self):
        sv = self.sv
        if self.pipeline:
            Lg, Rg = self._front(self.Ld, self.Rd)
            self.m._run(Lg, Rg, self.local, out=self.disp_d)
            self.xyz = sv.reprojectImageTo3D(self.disp_d, self.Q)
        else:
            self.m._run(self.Ld, self.Rd, self.local, out=self.disp_d)
        return sv.box_median_depth(self.disp_d, self.Q, self.bx, box_image=self.bi)

    def step_e2e(self):
        """Through the public API from pinned host buffers and back (copies inside the timed region)."""
        if not self.pipeline:   # streaming front end: copies of neighbouring steps overlap the matcher
            self.pipe.submit(self.Lh, self.Rh, self.disp_h, Q=self.Q, boxes=self.bx, box_image=self.bi, med_h=self.med_h)
            return
        sv = self.sv
        self.Le.copy_(self.Lh, non_blocking=True)
        self.Re.copy_(self.Rh, non_blocking=True)
        Lg, Rg = self._front(self.Le, self.Re)
        self.m._run(Lg, Rg, self.local, out=self.disp_d)
        xyz = sv.reprojectImageTo3D(self.disp_d, self.Q)
        med = sv.box_median_depth(self.disp_d, self.Q, self.bx, box_image=self.bi)
        self.disp_h.copy_(self.disp_d, non_blocking=True)
        self.xyz_h.copy_(xyz, non_blocking=True)
        self.med_h.copy_(med, non_blocking=True)

    def flush(self):
        if not self.pipeline:
            self.pipe.flush()

    def close(self):
        self.sv.release_workspaces()


def run_ours(args):
    import torch
    import torch.distributed as dist
    import stereo_vision_b200 as sv
    from stereo_vision_b200 import shard

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = CONFIGS[args.config]
    B = args.batch or cfg["batch"]
    Hc, Wc, Dc = cfg["H"], cfg["W"], cfg["D"]
    wl = Workload(args.config, B, rank, local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, collect=None, finish=None, sync=True):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier() if sync else torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            fn()
            if collect is not None:
                collect()
        if finish is not None:
            finish()                                    # the timed stream waits for the last device->host copies
        e1.record()
        barrier() if sync else torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        return shard.max_over_ranks(ms, dev) if sync else ms   # device time, max over ranks

    for _ in range(max(args.warmup, 3)):
        wl.step_resident()
    for _ in range(2):
        wl.step_e2e()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    # --- device-resident (value) with per-stage events on the launching stream
    sv.set_stage_timing(True, local)
    stage_acc = {}

    def collect():
        for k, v in sv.last_stage_ms(local).items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    timed(wl.step_resident, args.steps, collect)
    sv.set_stage_timing(False, local)
    ms_res_clean = timed(wl.step_resident, args.steps)     # without the per-step event sync
    ms_e2e = timed(wl.step_e2e, args.steps, finish=wl.flush)
    clocks = sampler.stop() if rank == 0 else None
    # parity of what was just timed: every frame of the last e2e step against cv2's hashes (rank 0 holds the golden seeds)
    parity = None
    if rank == 0:
        torch.cuda.synchronize()
        parity = check_parity(args.config, wl)
    if world > 1:
        dist.barrier()
    extra = {}
    if rank == 0 and args.config in ("c3", "c4") and not args.no_other:
        wl.close()
        del wl.Ld, wl.Rd, wl.disp_d
        torch.cuda.empty_cache()
        extra = other_configs(local, timed)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    frames = B * world * args.steps
    fps = frames / (ms_res_clean / 1e3)
    fps_e2e = frames / (ms_e2e / 1e3)
    g = wl.m.geometry(Wc)
    W1, Dp = g["width1"], g["Dp"]
    stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
    line = {
        "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_res_clean / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic (%d distinct seeded frame pairs per GPU)" % B,
        "gpix_disp_per_s": fps * Wc * Hc * Dc / 1e9,
        "config": {"workload": cfg["desc"], "frames_per_step_per_gpu": B, "parallelism": f"frame-pair sharding x{world}, no collective",
                   "l2": "per-step working set (C+S volumes, %.1f GB/GPU) >> 126 MB L2; no explicit flush" % (2 * 2.0 * W1 * Hc * Dp * B / 1e9)},
        "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": wl.h2d * world,                 # all ranks
                "d2h_bytes_per_step": wl.d2h * world, "gpix_disp_per_s": fps_e2e * Wc * Hc * Dc / 1e9},
        "gpu_launches": wl.launches * args.steps,
        "clocks": clocks,
        "stage_ms_per_step": stage_ms,
        "parity_checked": bool(parity and parity["ok"]), "parity_detail": parity,
    }
    line.update(rooflines(args.config, wl, stage_ms, W1, Dp, local))
    line.update(extra)
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        nthr = max(1, min(cores, 32))
        v2, kind, c2, d2 = cpu_reference_fps(cfg, nthr, form="pool")
        v1, _, c1, d1 = cpu_reference_fps(cfg, nthr, form="single")
        line["cpu_baseline"] = {"value": v2, "unit": "frames/s", "cores": c2, "kind": kind, "sample": d2,
                                "single_call": {"value": v1, "unit": "frames/s", "cores": c1, "sample": d1,
                                                "note": "BASELINE.md §3 row (i): cv2.setNumThreads(all), one compute() at a time"}}
    else:
        line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": 0, "kind": "skipped", "sample": ""}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0, help="frame pairs per step per GPU (default: the configuration's; c3/c4: 32)")
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS), help="BASELINE.json configuration to run as the main workload")
    ap.add_argument("--no-other", action="store_true", help="skip the other_configs / latency_ms block of the default run")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
