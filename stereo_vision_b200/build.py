"""Build libsvb.so (hand-written sm_100a CUDA kernels + C ABI) in-tree with nvcc.

    python -m stereo_vision_b200.build [--force] [--verbose]

The shared library lands at stereo_vision_b200/libsvb.so (git-ignored; travels to the GPU box with gpurun).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
SO = os.path.join(HERE, "libsvb.so")
SOURCES = ["svb_api.cu", "svb_rectify.cu", "svb_cost.cu", "svb_aggregate.cu", "svb_post.cu", "svb_depth.cu", "svb_bench.cu", "svb_sparse.cu"]
AGG_PARTS = 7
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-Xcompiler", "-fvisibility=default", "--threads", "2"]


def _deps_mtime() -> float:
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(os.path.dirname(HERE), "include", "svb.h"))
    return max(os.path.getmtime(h) for h in hdrs)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hm = _deps_mtime()
    todo, objs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        # the sweep kernel families of svb_aggregate.cu are separate objects (they dominate the build time)
        parts = [(f"_p{k}", [f"-DSVB_AGG_PART={k}"]) for k in range(AGG_PARTS)] if s == "svb_aggregate.cu" else [("", [])]
        for tag, defs in parts:
            obj = os.path.join(OBJ, s[:-3] + tag + ".o")
            objs.append(obj)
            if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hm):
                todo.append((src, obj, defs))
    todo.sort(key=lambda j: 0 if "svb_aggregate" in j[0] else 1)   # longest jobs first

    def cc(job):
        src, obj, defs = job
        cmd = [NVCC, *FLAGS, *defs, "-c", src, "-o", obj] + (["-Xptxas", "-v"] if verbose else [])
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(cc, todo))
    if todo or not os.path.exists(SO):
        cmd = [NVCC, "-shared", "-o", SO, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
