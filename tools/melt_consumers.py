#!/usr/bin/env python
"""tools/melt_consumers.py -- cfg4 through melt + libmltcuda.so: downloading null consumer vs the device-frame
cuda.null consumer, with and without the CPU filter on track 1.  Steady-state frames/s
(bench.melt_run: melt's -progress2 position, timestamped)."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

if __name__ == "__main__":
    threads = [int(t) for t in (sys.argv[1] if len(sys.argv) > 1 else "8").split(",")]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 12000
    profile = sys.argv[3] if len(sys.argv) > 3 else "uhd_2160p_60"
    base = tempfile.mkdtemp()
    noflip = [a for a in bench.MELT_CFG4 if a not in ("-attach", "mirror:flip")]
    for name, graph in (("cfg4", bench.MELT_CFG4), ("cfg4 without mirror:flip", noflip)):
        for consumer in ("null", "cuda.null"):
            for t in threads:
                r = bench.melt_run(n, t, True, base, 1024, graph=graph, consumer=consumer, profile=profile, extra={"MLT_CUDA_STATS": "1"})
                st = bench.parse_cuda_stats(r["stderr"])
                g = lambda k: st.get(k, {}).get("calls", 0)
                print(f"{name:26s} {profile} -consumer {consumer:9s} real_time=-{t:<2d}: steady {r['steady_fps']:8.1f} frames/s, whole run "
                      f"{r['whole_run_fps']:8.1f}, start-up {r['startup_s']:.2f} s, downloads {g('download + sync')}, uploads {g('upload')}, "
                      f"fused {g('fused pass (k_tractor launches)')}", flush=True)
