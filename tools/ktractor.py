#!/usr/bin/env python
"""tools/ktractor.py -- the fused kernel alone on cfg4 (the roofline leg of bench.py, nothing else): kernel-only
frames/s and fraction of the measured HBM peak.  Fast enough to iterate on, and the command ncu wraps.
  python tools/ktractor.py [--steps K] [--streams S] [--w W --h H] [--generic]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import bench  # noqa: E402
from mlt_b200 import Context, abi, workloads  # noqa: E402


class NoDist:
    rank, local_rank, world = 0, 0, 1

    def barrier(self):
        pass

    def reduce(self, x, op):
        return x


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--w", type=int, default=3840)
    ap.add_argument("--h", type=int, default=2160)
    ap.add_argument("--distinct", type=int, default=32)
    ap.add_argument("--repeat", type=int, default=8)
    ap.add_argument("--graph", action="store_true", help="replay the step from a CUDA graph instead of launching from python")
    ap.add_argument("--interp", default="bilinear", help="affine sampler: nearest | bilinear | bicubic (non-bilinear = the ONLY_BL=false instantiation)")
    a = ap.parse_args()
    lib = abi.load()
    workers = [Context(0) for _ in range(a.streams)]
    if a.interp != "bilinear":
        _ap = workloads.affine_params
        workloads.affine_params = lambda *x, **k: _ap(*x[:7], a.interp)
    S = bench.cfg4_setup(lib, workers[0], workloads.interleaved_frames(0, 1, a.distinct), a.w, a.h)
    full, fused_only = bench.cfg4_steps(lib, S, workers, a.repeat)
    timed = bench.Timed(lib, workers)
    fused_only()
    if a.graph:
        ms, launches = timed.run_graph(timed.capture(fused_only), a.steps, NoDist())
    else:
        ms, launches = timed.run(fused_only, a.steps, NoDist())
    cov = [bench.covered_pixels(f["affine"], a.w, a.h) for f in S["frames"]]
    alg = float(np.mean([a.w * a.h * 5 + S["b1b"] + S["b3b"] + 4 * c for c in cov]))
    us = ms * 1e3 / launches
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)
    print(f"k_tractor {a.w}x{a.h} {a.interp}{' generic' if os.environ.get('MLTCUDA_TRACTOR_GENERIC') else ''}: {launches} launches, {us:.2f} us/launch, {1e6 / us:.0f} frames/s, "
          f"{alg / us / 1e3:.1f} GB/s algorithmic = {alg / us / 1e3 / peak:.4f} of measured {peak} GB/s")
