#!/usr/bin/env python
"""tools/melt_bench.py -- cfg4 through the reference's own melt, stock services vs the CUDA module.
Prints frames/s for each (wall clock over the whole melt run, -consumer null real_time=-N)."""
import argparse
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mltenv  # noqa: E402

CFG4 = ["noise:", "out={last}", "-track", "noise:", "out={last}", "-attach", "mirror:flip", "-transition", "composite",
        "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-track", "colour:0x3060c0a0", "out={last}", "-transition",
        "affine", "a_track=0", "b_track=2", "rect=0=20%/20%:50%x50%:90%; 599=30%/25%:60%x60%:70%", "rotate_x=0.5", "-track",
        "noise:", "out={last}", "-transition", "composite", "a_track=0", "b_track=3", "geometry=55%/55%:40%x40%:100%"]


def run(cuda, frames, threads, profile, base):
    e = mltenv.env(base, cuda=True, override=cuda)
    args = [a.format(last=frames - 1) for a in CFG4]
    cmd = [mltenv.MELT, "-profile", profile] + args + ["-consumer", "null", f"real_time=-{threads}", "-silent"]
    t0 = time.perf_counter()
    r = subprocess.run(cmd, env=e, capture_output=True, text=True)
    dt = time.perf_counter() - t0
    if r.returncode:
        print(r.stderr[-2000:])
    return frames / dt, dt


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=12000)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--profile", default="uhd_2160p_60")
    a = ap.parse_args()
    base = tempfile.mkdtemp()
    # Two run lengths per arm: the difference removes process start-up (CUDA initialisation, one
    # context + pinned staging per worker thread, first-use kernel loading: 1.5-3 s for the module,
    # 0.1 s for the stock services) and gives the steady-state rate a long timeline sees.
    short = max(8, a.frames // 12)
    for cuda in (False, True):
        n = a.frames if cuda else max(16, min(400, a.frames // 8))  # the stock arm is ~100x slower: fewer frames
        s0 = short if cuda else max(8, n // 6)
        _, t1 = run(cuda, s0, a.threads, a.profile, base)
        fps, t2 = run(cuda, n, a.threads, a.profile, base)
        a_frames, short_n = n, s0
        steady = (a_frames - short_n) / max(t2 - t1, 1e-9)
        print(f"{'cuda module' if cuda else 'stock (CPU)'}: {fps:.2f} frames/s over the whole run ({a_frames} frames, {t2:.2f} s), "
              f"steady state {steady:.1f} frames/s, start-up ~{t1 - short_n / steady:.2f} s "
              f"(real_time=-{a.threads}, {a.profile})", flush=True)
