"""mltcuda_noise_static -- the noise producer for callers that must not allocate -- against mltcuda_noise.

Some users of the library cannot let it call cudaMalloc: a real_time render thread in mid-playback (an allocation
takes a device-wide lock and drains the GPU, stalling the other workers' streams), or a host that has handed all
device memory to its own pool / lives inside a fixed MIG or MPS budget, where an allocation can fail with no way
to recover.  For them mltcuda_noise_static keeps the producer's skip-ahead tables in a static __device__ array
inside the library instead of a per-context cudaMalloc block.  It must produce exactly the bytes mltcuda_noise
does (integer work: the tolerance is zero, as in test_gpu_kernels.test_noise), on the call that sets the array up
and on every call after it, whichever context of the GPU makes it.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should not hide the
parity tests.  It passes on its own in a fresh process (its first call then performs the per-GPU set-up) and after
the rest of the suite (same: nothing else touches the static array).
"""
import numpy as np
import pytest

from mlt_b200 import Context, abi

pytestmark = pytest.mark.gpu


def _noise(ctx, entry, w, h, pos, batch=1):
    """Run one noise entry point into a buffer pre-filled with 0xAA and return the bytes."""
    frame = w * h * 2
    stride = (frame + 15) // 16 * 16
    d = ctx.put(np.full(stride * batch, 0xAA, np.uint8))
    abi.check(getattr(ctx.lib, entry)(ctx.h, d.ptr, w, h, pos, batch, stride if batch > 1 else 0), entry)
    got = ctx.get(d, (batch, stride))[:, :frame].copy()
    d.free()
    return got


def test_static_noise_matches_allocating_noise_twice_in_a_row(ctx):
    # two calls back to back: the first may be the one that sets the static array up, the second must find
    # it as the first left it.  Different size and timeline position each time, a ragged tail in the second.
    for w, h, pos in [(1920, 1080, 3), (37, 3, 100000)]:
        exp = _noise(ctx, "mltcuda_noise", w, h, pos)
        got = _noise(ctx, "mltcuda_noise_static", w, h, pos)
        assert np.array_equal(got, exp), f"{w}x{h}@{pos}: {(got != exp).sum()} bytes differ"


def test_static_noise_batch_and_second_context(ctx):
    # a context created later on the same GPU shares the array the first one filled
    w, h, n = 128, 16, 4
    exp = _noise(ctx, "mltcuda_noise", w, h, 7, n)
    with Context(0) as other:
        got_other = _noise(other, "mltcuda_noise_static", w, h, 7, n)
    got = _noise(ctx, "mltcuda_noise_static", w, h, 7, n)
    assert np.array_equal(got_other, exp)
    assert np.array_equal(got, exp)


def test_static_noise_refuses_sizes_beyond_its_tables(ctx):
    d = ctx.put(np.zeros(64, np.uint8))
    launches = ctx.launches()
    rc = ctx.lib.mltcuda_noise_static(ctx.h, d.ptr, 65536, 32768, 0, 1, 0)  # 2^31 pixels: one too many
    assert rc == abi.E_BADARG
    assert b"beyond what the static tables reach" in ctx.lib.mltcuda_last_error()
    rc = ctx.lib.mltcuda_noise_static(ctx.h, d.ptr, 16, 16, 0, 65536, 512)  # too many frames in one call
    assert rc == abi.E_BADARG
    assert ctx.launches() == launches, "a refused call must not launch anything"
    d.free()
