#!/usr/bin/env python
"""tools/cfg5.py -- BASELINE.json config 5: the 10-minute uhd_2160p_60 4-track timeline (36 000 frames) rendered
frame-parallel through ONE melt process whose real_time worker threads map onto all visible GPUs (worker k -> GPU
k mod G, module/cuda_image.c), against the stock services on the host cores (real_time=-nproc, bounded sample).
  python tools/cfg5.py [--frames 36000] [--threads 32] [--ref-frames 1600] [--consumer null]"""
import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mlt_b200 import abi  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=36000)
    ap.add_argument("--threads", default="")
    ap.add_argument("--ref-frames", type=int, default=1600)
    ap.add_argument("--consumers", default="null,cuda.null")
    a = ap.parse_args()
    g = abi.load().mltcuda_device_count()
    cores = os.cpu_count() or 1
    base = tempfile.mkdtemp()
    # MLT producers default to length 15000 (out is clamped to it): a 36 000-frame timeline needs length= on every track
    graph = []
    for tok in bench.MELT_CFG4:
        graph.append(tok)
        if tok == "out={last}":
            graph.append(f"length={a.frames}")
    for t in [int(x) for x in (a.threads or f"{4 * g}").split(",")]:
        for consumer in a.consumers.split(","):
            r = bench.melt_run(a.frames, t, True, base, 2048, tail=256, consumer=consumer, graph=graph, extra={"MLT_CUDA_STATS": "1"})
            st = bench.parse_cuda_stats(r["stderr"])
            print(json.dumps({"config": "cfg5_uhd_2160p_60_4track_10min", "impl": "cuda module, one melt process", "gpus": g,
                              "consumer": consumer, "real_time": -t, "frames": a.frames, "steady_fps": round(r["steady_fps"], 1),
                              "whole_run_fps": round(r["whole_run_fps"], 1), "wall_s": round(r["total_s"], 2), "window": r["window"],
                              "fused_launches": st.get("fused pass (k_tractor launches)", {}).get("calls"),
                              "downloaded_GB": round(st.get("download + sync", {}).get("bytes", 0) / 1e9, 1),
                              "uploaded_GB": round(st.get("upload", {}).get("bytes", 0) / 1e9, 1)}), flush=True)
    if a.ref_frames > 0:
        r = bench.melt_run(a.ref_frames, cores, False, base, 128, tail=64)
        print(json.dumps({"config": "cfg5_uhd_2160p_60_4track_10min", "impl": "reference (stock services)", "cores": cores,
                          "real_time": -cores, "frames": a.ref_frames, "note": "bounded sample of the timeline",
                          "steady_fps": round(r["steady_fps"], 2), "whole_run_fps": round(r["whole_run_fps"], 2),
                          "wall_s": round(r["total_s"], 1), "projected_wall_s_36000": round(36000 / r["steady_fps"], 0)}), flush=True)
