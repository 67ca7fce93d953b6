#!/usr/bin/env python
"""tools/melt_scan.py -- where does the melt-level (module) frame rate go?  Steady-state frames/s of the
CUDA module for a few graphs and worker-thread counts (difference of two run lengths, like melt_bench)."""
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import mltenv  # noqa: E402
from melt_bench import CFG4  # noqa: E402

GRAPHS = {
    "noise only": ["noise:", "out={last}"],
    "noise+composite": ["noise:", "out={last}", "-track", "noise:", "out={last}", "-transition", "composite", "a_track=0",
                        "b_track=1", "geometry=5%/5%:40%x40%:80%"],
    "cfg4": CFG4,
}


def run(graph, frames, threads, base, extra=()):
    e = mltenv.env(base, cuda=True, override=True)
    if os.environ.get("MLT_CUDA_STATS"):
        e["MLT_CUDA_STATS"] = "1"
    args = [a.format(last=frames - 1) for a in graph]
    rt = f"real_time=-{threads}" if threads else "real_time=0"  # 0: frames are built but never rendered
    cmd = [mltenv.MELT, "-profile", "uhd_2160p_60"] + args + ["-consumer", "null", rt, "-silent"] + list(extra)
    t0 = time.perf_counter()
    r = subprocess.run(cmd, env=e, capture_output=True, text=True)
    dt = time.perf_counter() - t0
    if r.returncode:
        print(r.stderr[-1000:])
    for ln in r.stderr.splitlines():
        if ln.startswith("[cuda stats]"):
            print("    " + ln)
    return dt


if __name__ == "__main__":
    base = tempfile.mkdtemp()
    n0, n1 = 16, 4500
    for name, g in GRAPHS.items():
        if name == "noise only":
            continue
        for th in (16,):
            t0 = run(g, n0, th, base)
            t1 = run(g, n1, th, base)
            print(f"{name:18s} real_time=-{th:<3d} steady {(n1 - n0) / max(t1 - t0, 1e-9):8.1f} frames/s  (start-up ~{t0 - n0 * (t1 - t0) / (n1 - n0):.2f} s)", flush=True)
