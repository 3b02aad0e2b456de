#!/usr/bin/env python
"""bench.py — 1080p D=256 SGBM throughput (frames/s, Gpix*disp/s) on N B200s vs the cv2 CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]

Workload (BASELINE.json configs[2]/[3], "c3"/"c4"): synthetic rectified 1920x1080 gray pairs, numDisparities=256,
blockSize=5, MODE_HH (8 paths), LR check, uniqueness, 3x3 median, speckle filter, then per-YOLO-box median depth
(8 synthetic boxes per frame).  One "step" = one batch of B frame pairs per GPU through that hot path.  Frame
pairs are independent, so ranks shard the batch with no data-path collective (weak scaling; NCCL is used only
for the barrier / max-over-ranks timing).

Prints ONE JSON line (rank 0).  See DESIGN.md §measurement for how each field is obtained.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, D, BS, MODE = 1080, 1920, 256, 5, 1
NBOX = 8
WORKLOAD = "c3/c4: 1920x1080 pairs, numDisparities=256, blockSize=5, MODE_HH 8-path, LR check + speckle, 8 box medians/frame"


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _ncu_traffic(B: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed ncu --set full capture
    (profiles/r01_traffic.json, per launch at the batch it was captured with; scaled linearly to this batch)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            t = json.load(f)["k_sweep_pass2"]
        return float(t["dram_bytes"]) * B / t["batch"], "%s (B=%d, scaled x%g)" % (t["source"], t["batch"], B / t["batch"])
    except Exception:
        return None, "no capture committed"


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


def make_batch(B: int, rank: int):
    from stereo_vision_b200 import synth
    # a few distinct seeded frames per rank, tiled to B (content does not change the arithmetic done)
    nuniq = min(B, 4)
    pairs = [synth.stereo_pair(H, W, D, seed=1000 + rank * 64 + i) for i in range(nuniq)]
    L = np.stack([pairs[i % nuniq][0] for i in range(B)])
    R = np.stack([pairs[i % nuniq][1] for i in range(B)])
    boxes = np.concatenate([synth.boxes_for_frame(H, W, D, NBOX, seed=rank * 64 + i) for i in range(B)])
    box_image = np.repeat(np.arange(B, dtype=np.int32), NBOX)
    K1, _, K2, _, Rm, T = synth.synthetic_rig(W, H)
    Q = synth.simple_rectify(K1, K2, Rm, T, W, H)[4]
    return L, R, boxes, box_image, Q


def cpu_reference_fps(n_threads: int, frames_per_thread: int = 1):
    """Times the reference's own CPU implementation of the path: cv2.StereoSGBM (all host threads via a pool over
    frame pairs, cv2.setNumThreads(1) inside: cv2's SGBM core is single-threaded) or, when cv2 is
    not importable, the C oracle port.  Returns (frames/s, kind, cores, sample description)."""
    from concurrent.futures import ThreadPoolExecutor
    from stereo_vision_b200 import synth
    kw = synth.sgbm_kwargs(D, BS, MODE)
    L, R = synth.stereo_pair(H, W, D, seed=1000)
    try:
        import cv2
        cv2.setNumThreads(1)
        kind = "reference"

        K1, _, K2, _, Rm, T = synth.synthetic_rig(W, H)
        Q = synth.simple_rectify(K1, K2, Rm, T, W, H)[4]
        boxes = synth.boxes_for_frame(H, W, D, NBOX, seed=0)

        def one(_):
            m = cv2.StereoSGBM_create(**kw)
            for _ in range(frames_per_thread):
                d = m.compute(L, R)
                z = cv2.reprojectImageTo3D(d.astype(np.float32) / 16, Q)[..., 2]
                for (x1, y1, x2, y2) in boxes:   # per-box median depth, as the GPU arm computes
                    sl = (slice(int(np.floor(y1)), int(np.ceil(y2))), slice(int(np.floor(x1)), int(np.ceil(x2))))
                    v = z[sl][(d[sl] != -16) & np.isfinite(z[sl])]
                    _ = np.median(v) if v.size else np.nan
        desc = f"cv2 {cv2.__version__} StereoSGBM.compute, {n_threads} threads x {frames_per_thread} 1080p D=256 MODE_HH pairs"
        t0 = time.perf_counter()
        with ThreadPoolExecutor(n_threads) as ex:
            list(ex.map(one, range(n_threads)))
        dt = time.perf_counter() - t0
        nfr = n_threads * frames_per_thread
    except ImportError:
        import oracle
        kind = "port"
        oracle.set_num_threads(n_threads)
        nfr = 1
        desc = f"oracle/sv_oracle.c (OpenMP, {n_threads} threads), 1 x 1080p D=256 MODE_HH pair"
        t0 = time.perf_counter()
        oracle.sgbm_compute(L, R, **kw)
        dt = time.perf_counter() - t0
    return nfr / dt, kind, n_threads, desc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    nthr = max(1, min(cores, 32))
    vals = []
    for i in range(args.warmup + args.steps):
        fps, kind, c, desc = cpu_reference_fps(nthr)
        if i >= args.warmup:
            vals.append(fps)
    v = float(np.mean(vals))
    line = {"impl": "reference", "metric": "frames_per_s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * nthr / v, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
            "gpix_disp_per_s": v * W * H * D / 1e9,
            "config": {"workload": WORKLOAD, "frames_per_step": nthr, "host_cores": cores},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": c, "kind": kind, "sample": desc},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    import stereo_vision_b200 as sv
    from stereo_vision_b200 import shard, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    L, R, boxes, box_image, Q = make_batch(B, rank)
    kw = synth.sgbm_kwargs(D, BS, MODE)
    m = sv.StereoSGBM_create(**kw)
    Ld, Rd = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
    bx, bi = torch.from_numpy(boxes).to(dev), torch.from_numpy(box_image).to(dev)
    # pinned host staging for the end-to-end arm
    Lh, Rh = torch.from_numpy(L).pin_memory(), torch.from_numpy(R).pin_memory()
    disp_h = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
    med_h = torch.empty(B * NBOX, dtype=torch.float32).pin_memory()
    disp_d = torch.empty((B, H, W), dtype=torch.int16, device=dev)

    def step_resident():
        m._run(Ld, Rd, local, out=disp_d)
        return sv.box_median_depth(disp_d, Q, bx, box_image=bi)

    # end to end through the public streaming front end: every step copies its inputs from pinned host memory and its
    # disparities + box medians back; copies of neighbouring steps overlap the matcher (three streams, two slots)
    pipe = sv.HostPipeline(m, device=local)

    def step_e2e():
        pipe.submit(Lh, Rh, disp_h, Q=Q, boxes=bx, box_image=bi, med_h=med_h)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, collect=None, finish=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
            if collect is not None:
                collect()
        if finish is not None:
            finish()                                    # the timed stream waits for the last device->host copies
        e1.record()
        barrier()
        return shard.max_over_ranks(e0.elapsed_time(e1), dev)   # device time, max over ranks

    for _ in range(max(args.warmup, 3)):
        step_resident()
    for _ in range(2):
        step_e2e()
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
    ms_res = timed(step_resident, args.steps, collect)
    sv.set_stage_timing(False, local)
    ms_res_clean = timed(step_resident, args.steps)     # without the per-step event sync
    ms_e2e = timed(step_e2e, args.steps, finish=pipe.flush)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        dist.destroy_process_group()
    if rank != 0:
        return
    frames = B * world * args.steps
    fps = frames / (ms_res_clean / 1e3)
    fps_e2e = frames / (ms_e2e / 1e3)
    g = m.geometry(W)
    W1, Dp = g["width1"], g["Dp"]
    stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
    # dominant kernel: the pass-2 wavefront sweep (k_sweep<..,NV=3,PASS2>): reads C and S once per launch
    dom = "aggregate_pass2_select"
    alg_bytes = 4.0 * W1 * H * Dp * B                      # 2 B of C + 2 B of S per (pixel, disparity)
    peak, peak_src = _peaks()
    ach = alg_bytes / (stage_ms[dom] / 1e3) / 1e9
    # 9 16-bit ops per (px, d, path); 4 paths per pass, 8 per frame pair (MODE_HH)
    ops16_pass = 9.0 * W1 * H * D * 4 * B
    agg_ms = stage_ms["aggregate_pass1"] + stage_ms[dom]
    alu = sv.packed_alu_peak(local)
    traffic, traffic_src = _ncu_traffic(B)
    cores = os.cpu_count() or 1
    cpu_v, cpu_kind, cpu_c, cpu_desc = cpu_reference_fps(max(1, min(cores, 32))) if not args.no_cpu else (None, "skipped", 0, "")
    line = {
        "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_res_clean / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
        "gpix_disp_per_s": fps * W * H * D / 1e9,
        "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "parallelism": f"frame-pair sharding x{world}, no collective",
                   "l2": "per-step working set (C+S volumes, %.1f GB/GPU) >> 126 MB L2; no explicit flush" % (2 * 2.0 * W1 * H * Dp * B / 1e9)},
        "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": int(2 * B * H * W) * world,                 # all ranks
                "d2h_bytes_per_step": int(B * H * W * 2 + B * NBOX * 4) * world, "gpix_disp_per_s": fps_e2e * W * H * D / 1e9},
        "gpu_launches": 11 * args.steps,   # planes, cost, sweep x2, lrcheck, median3, ccl x4, box_median
        "clocks": clocks,
        "stage_ms_per_step": stage_ms,
        "roofline": {"kernel": "k_sweep<NR=4,G=32,NGRP=15,NV=3,PASS2> (aggregation pass 2 + selection)", "bound": "hbm",
                     "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                     "traffic_source": traffic_src,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                     "share_of_step": stage_ms[dom] / sum(stage_ms.values()),
                     "note": "this kernel is packed-int ALU-bound by design (north_star); see alu_roofline"},
        # north_star's aggregation roofline: algorithmic 16-bit ops (9 per px*d*path) against the packed-int ALU peak
        # measured live by svb_measure_packed_alu_peak (VIADDMNMX/VIMNMX.S16x2 chains on all SMs, x2 lanes)
        "alu_roofline": {"stage": "aggregation (both sweeps)", "achieved_ops16_per_s": 2 * ops16_pass / (agg_ms / 1e3),
                         "peak_ops16_per_s": alu["ops16_per_s"], "frac": 2 * ops16_pass / (agg_ms / 1e3) / alu["ops16_per_s"],
                         "pass2_only_frac": ops16_pass / (stage_ms[dom] / 1e3) / alu["ops16_per_s"],
                         "peak_source": "measured live (svb_measure_packed_alu_peak)"},
        "cpu_baseline": {"value": cpu_v, "unit": "frames/s", "cores": cpu_c, "kind": cpu_kind, "sample": cpu_desc},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32, help="frame pairs per step per GPU (c4: 256 pairs over 8 GPUs)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
