#!/usr/bin/env python
"""tools/pcie_ceiling.py -- what the box's PCIe / host-memory path gives N GPUs at once: plain pinned
cudaMemcpyAsync loops through the C-ABI (mltcuda_upload / mltcuda_download), one process per GPU, all started
together.  This is the ceiling of every end-to-end number that moves frames to host memory.
  python tools/pcie_ceiling.py [--gpus 1,2,4,8] [--mb 16.6] [--seconds 2] [--numa]"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(dev, nbytes, seconds, mode, barrier, q, pin_cpus):
    from mlt_b200 import Context, abi
    if pin_cpus:
        try:
            os.sched_setaffinity(0, pin_cpus)
        except OSError:
            pass
    lib = abi.load()
    ctx, ctx2 = Context(dev), Context(dev)
    slots = 4
    h = [lib.mltcuda_host_alloc(nbytes) for _ in range(slots)]
    d = [ctx.alloc(nbytes) for _ in range(slots)]
    for i in range(slots):
        lib.mltcuda_upload(ctx.h, d[i].ptr, h[i], nbytes)
    ctx.sync()
    barrier.wait()
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        for i in range(slots):
            if mode in ("h2d", "both"):
                lib.mltcuda_upload(ctx.h, d[i].ptr, h[i], nbytes)
            if mode in ("d2h", "both"):
                lib.mltcuda_download(ctx2.h, h[(i + 2) % slots], d[(i + 2) % slots].ptr, nbytes)
        ctx.sync()
        ctx2.sync()
        n += slots
    dt = time.perf_counter() - t0
    q.put((dev, n * nbytes / dt / 1e9))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1")
    ap.add_argument("--mb", type=float, default=16.6)
    ap.add_argument("--seconds", type=float, default=2.0)
    ap.add_argument("--numa", action="store_true", help="pin each process to the CPUs of its GPU's NUMA node (if sysfs tells)")
    a = ap.parse_args()
    from mlt_b200 import abi
    ndev = abi.load().mltcuda_device_count()
    nbytes = int(a.mb * 1e6) // 16 * 16
    mp.set_start_method("spawn")
    for g in [int(x) for x in a.gpus.split(",")]:
        if g > ndev:
            continue
        for mode in ("h2d", "d2h", "both"):
            barrier, q = mp.Barrier(g), mp.Queue()
            ncpu = os.cpu_count() or 1
            ps = []
            for dev in range(g):
                cpus = set(range(dev * ncpu // g, (dev + 1) * ncpu // g)) if a.numa else None
                p = mp.Process(target=worker, args=(dev, nbytes, a.seconds, mode, barrier, q, cpus))
                p.start()
                ps.append(p)
            res = sorted(q.get() for _ in range(g))
            for p in ps:
                p.join()
            per = [round(r[1], 1) for r in res]
            print(json.dumps({"gpus": g, "mode": mode, "frame_mb": a.mb, "aggregate_gbs_per_direction": round(sum(per), 1),
                              "per_gpu_gbs": per, "pinned_cpus": bool(a.numa)}), flush=True)
