"""GPU parity for the whole cfg4 pass: fused tractor kernel == one-launch-per-op == CPU oracle chain,
plus size-independent properties at the full BASELINE size (uhd_2160p)."""
import numpy as np
import pytest

import oracle
import pipeline
from mlt_b200 import abi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("W,H,frame", [(640, 360, 0), (640, 360, 123), (1920, 1080, 37)])
def test_cfg4_fused_unfused_oracle(ctx, W, H, frame):
    inp = pipeline.inputs(frame, W, H)
    exp = pipeline.oracle_chain(inp)
    unf = pipeline.gpu_unfused(ctx, inp)
    fus = pipeline.gpu_fused(ctx, inp)
    # fused and unfused run identical device arithmetic: byte-identical
    assert np.array_equal(fus["out"], unf["out"])
    assert np.array_equal(fus["alpha"], unf["alpha"])
    # vs the oracle (pinned to the compiled reference by tests/test_golden.py): the kernels follow the Release
    # build's operation order, so the whole chain is byte-identical, float affine included
    assert np.array_equal(fus["out"], exp["out"])
    assert np.array_equal(fus["alpha"], exp["alpha"])


@pytest.mark.parametrize("interp", ["nearest", "bicubic"])
def test_cfg4_other_samplers_fused_unfused_oracle(ctx, interp):
    """consumer.rescale=nearest takes k_tractor's nearest-only instantiation (transposed pass), bicubic the run-time one."""
    inp = pipeline.inputs(41, 1920, 1080, interp)
    exp = pipeline.oracle_chain(inp)
    unf = pipeline.gpu_unfused(ctx, inp)
    fus = pipeline.gpu_fused(ctx, inp)
    assert np.array_equal(fus["out"], unf["out"]) and np.array_equal(fus["alpha"], unf["alpha"])
    diff = np.abs(fus["out"].astype(int) - exp["out"].astype(int))
    assert diff.max() <= 2 and (diff > 0).mean() < 1e-4, (int(diff.max()), float((diff > 0).mean()))


def test_cfg4_uhd_properties(ctx):
    """Full-size (3840x2160) checks that need no CPU pass over the frame: idempotent inputs,
    untouched regions stay equal to the bottom track, fused == unfused."""
    W, H, frame = 3840, 2160, 300
    inp = pipeline.inputs(frame, W, H)
    unf = pipeline.gpu_unfused(ctx, inp)
    fus = pipeline.gpu_fused(ctx, inp)
    assert np.array_equal(fus["out"], unf["out"])
    assert np.array_equal(fus["alpha"], unf["alpha"])
    out = fus["out"].reshape(H, W, 2)
    a = inp["a"].reshape(H, W, 2)
    # top-left 5% border is outside every layer: yuv422 -> rgba -> yuv422 of noise (U=V=128) ...
    _, rgba, _ = oracle.convert(3, 2, W, 8, inp["a"][:W * 8 * 2])
    _, back, _ = oracle.convert(2, 3, W, 8, rgba, want_alpha=True)
    assert np.array_equal(out[:8].reshape(-1), back)
    # the second composite (100 %, opaque B, alpha_a present => SSE2 case 4) is a blit of track 3
    g2, b2w, b2h = inp["f"]["comp2"]
    x, y = int(round(g2["x"])), int(round(g2["y"]))
    assert np.array_equal(out[y:y + b2h, x:x + b2w].reshape(-1), inp["b3"])
    assert (fus["alpha"].reshape(H, W)[y:y + b2h, x:x + b2w] == 255).all()
    assert not np.array_equal(out, a)


def test_tractor_batch_equals_single(ctx):
    W, H, n = 640, 360, 3
    frames = [pipeline.inputs(5, W, H) for _ in range(n)]
    rng = np.random.default_rng(0)
    for i, f in enumerate(frames):
        f["a"] = rng.integers(16, 240, W * H * 2).astype(np.uint8)
    single = [pipeline.gpu_fused(ctx, f)["out"] for f in frames]
    d_in = ctx.put(np.concatenate([f["a"] for f in frames]))
    d_b1, d_bc, d_b3 = ctx.put(frames[0]["b1"]), ctx.put(frames[0]["bc"]), ctx.put(frames[0]["b3"])
    d_out, d_al = ctx.alloc(n * W * H * 2), ctx.alloc(n * W * H)
    bs = {"in": W * H * 2, "out": W * H * 2, "out_alpha": W * H}
    ctx.call("mltcuda_tractor", pipeline.tractor_desc(ctx, frames[0]["f"], d_in.ptr, d_out.ptr, d_al.ptr, d_b1.ptr,
                                                      d_bc.ptr, d_b3.ptr, batch=n, bs=bs))
    got = ctx.get(d_out, (n, W * H * 2))
    for i in range(n):
        assert np.array_equal(got[i], single[i])


def test_tractor_rejects_bad_lists(ctx):
    import ctypes as C
    d = ctx.put(np.zeros(64 * 8 * 4, np.uint8))
    ops = (abi.TractorOp * 1)()
    ops[0].kind, ops[0].to_format = abi.TOP_CONVERT, 3  # yuv422 -> yuv422
    desc = abi.TractorDesc(width=64, height=8, in_format=3, out_format=3, in_=d.ptr, out=d.ptr, n_ops=1, ops=ops)
    assert ctx.lib.mltcuda_tractor(ctx.h, C.byref(desc)) == abi.E_BADARG
    desc = abi.TractorDesc(width=60, height=8, in_format=3, out_format=3, in_=d.ptr, out=d.ptr, n_ops=0)
    assert ctx.lib.mltcuda_tractor(ctx.h, C.byref(desc)) == abi.E_UNSUPPORTED
