"""N>1 host logic on CPU: two gloo ranks shard the timeline exactly as bench.py does under torchrun
(frame-parallel, no data-path collective), render their frames with the CPU oracle at a small size,
and the gathered result must equal the single-process rendering, frame for frame."""
import hashlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _digest(frame, W, H):
    import pipeline
    res = pipeline.oracle_chain(pipeline.inputs(frame, W, H))
    return hashlib.sha1(res["out"].tobytes() + res["alpha"].tobytes()).hexdigest()


def _worker(rank, world, port, F, W, H, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist
    import bench
    from mlt_b200 import workloads
    d = bench.Dist()  # the class bench.py uses: gloo barrier + reductions
    assert (d.rank, d.world) == (rank, world)
    frames = workloads.interleaved_frames(d.rank, d.world, F)  # what bench.py deals to each rank
    mine = [(f, _digest(f, W, H)) for f in frames]
    d.barrier()
    ms_max = d.reduce(10.0 + rank, "MAX")   # timing: max over ranks
    total = d.reduce(len(frames), "SUM")    # units: summed over ranks
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    q.put((rank, ms_max, total, gathered))
    d.close()


def test_two_gloo_ranks_shard_frames_like_bench():
    import torch.multiprocessing as mp
    world, F, W, H = 2, 3, 128, 72
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, F, W, H, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from mlt_b200 import workloads
    want = sorted(f for r in range(world) for f in workloads.interleaved_frames(r, world, F))
    assert len(set(want)) == world * F  # disjoint
    expected = {f: _digest(f, W, H) for f in want}
    for rank, ms_max, total, gathered in results:
        assert ms_max == 11.0 and total == world * F
        got = dict(x for part in gathered for x in part)
        assert sorted(got) == want  # disjoint and complete
        assert got == expected


def test_shard_frames_partitions_every_step():
    from mlt_b200 import workloads
    for world in (1, 2, 4, 8):
        for step in (0, 3):
            seen = []
            for r in range(world):
                seen += workloads.shard_frames(r, world, 32, step)
            assert seen == list(range(step * world * 32, (step + 1) * world * 32))


def test_interleaved_frames_give_every_rank_the_same_window():
    """bench.py's dealing: disjoint across ranks, and every rank (at any world size up to 8) walks the same
    32 x 8-frame window of the timeline with stride 8, so per-GPU work does not depend on the rank or on N."""
    from mlt_b200 import workloads
    for world in (1, 2, 4, 8, 16):
        stride = max(world, 8)
        allf = [workloads.interleaved_frames(r, world, 32) for r in range(world)]
        flat = [f for fr in allf for f in fr]
        assert len(set(flat)) == len(flat)
        for r, fr in enumerate(allf):
            assert fr == [i * stride + r for i in range(32)]
