"""Summarise an .ncu-rep (read here on the CPU box): per-kernel time, DRAM traffic, issue/pipe utilisation, top stalls.

    python tools/ncu_summary.py <report.ncu-rep> [--json profiles/r02_counters.json --batch B --source <name>]

With --json the per-kernel DRAM bytes and hardware pipe counters bench.py quotes (k_cost_fused, k_sweep pass 1 / 2) are
written to that file."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
def col(name):
    return hdr.index(name) if name in hdr else None
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_active.avg", "smsp__average_warp_latency_issue_stalled_barrier_per_warp_active.pct"]
stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
for r in data:
    print("==", r[col("Kernel Name")][:90])
    for w in want:
        c = col(w)
        if c is not None:
            print("   %-72s %s %s" % (w, r[c], units[c]))
    st = sorted(((float(r[hdr.index(s)] or 0), s) for s in stall), reverse=True)[:6]
    for v, s in st:
        print("   stall %-40s %.2f" % (s.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))

if "--json" in sys.argv:
    import json
    arg = lambda k, d=None: sys.argv[sys.argv.index(k) + 1] if k in sys.argv else d
    out_path, batch, source = arg("--json"), int(arg("--batch", "0")), arg("--source", rep)
    res = {}
    g = lambda r, n: float(r[col(n)].replace(",", "")) if col(n) is not None and r[col(n)] else None
    unit = lambda n: units[col(n)] if col(n) is not None else ""
    def to_bytes(v, u):
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
    for r in data:
        name = r[col("Kernel Name")]
        key = "k_cost_fused" if "k_cost_fused" in name else ("k_sweep_pass2" if "k_sweep" in name and ", 1, " in name.split("k_sweep")[1][:40] and "true" in name else None)
        if "k_sweep" in name:
            # template args: <NR, G, NGRP, NV, PASS2, SAFE, PAD, PFD>
            targs = name.split("<")[1].split(">")[0].split(",")
            pass2 = targs[4].strip() in ("1", "true", "(bool)1")
            key = "k_sweep_pass2" if pass2 else "k_sweep_pass1"
        if key is None:
            continue
        rd = to_bytes(g(r, "dram__bytes_read.sum"), unit("dram__bytes_read.sum"))
        wr = to_bytes(g(r, "dram__bytes_write.sum"), unit("dram__bytes_write.sum"))
        ms = g(r, "gpu__time_duration.sum")
        u = unit("gpu__time_duration.sum")
        ms = ms * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "nsecond": 1e-6, "s": 1e3, "second": 1e3}.get(u, 1)
        res[key] = {"kernel": name[:120], "batch": batch, "source": source, "time_ms_under_ncu": ms, "dram_bytes_read": rd, "dram_bytes_write": wr,
                    "dram_bytes": rd + wr, "registers": g(r, "launch__registers_per_thread"),
                    "pipe_alu_pct": g(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                    "pipe_fma_pct": g(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                    "issue_active_pct": g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "warps_active_pct": g(r, "sm__warps_active.avg.pct_of_peak_sustained_active")}
    json.dump(res, open(out_path, "w"), indent=1)
    print("wrote", out_path, list(res))
