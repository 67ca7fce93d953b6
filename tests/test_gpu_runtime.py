"""C-ABI runtime pieces the MLT module leans on: staged uploads from pageable memory, the pinned
host pool, cross-context ordering (include/mltcuda.h)."""
import ctypes as C

import numpy as np
import pytest

from mlt_b200 import Context, abi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nbytes", [1, 4095, 4 << 20, (4 << 20) + 1, (9 << 20) + 3, 33 << 20])
def test_upload_staged_roundtrip(ctx, nbytes):
    """Any size, incl. 0 / 1 / 2+ full 4 MiB staging slots plus a tail, arrives intact; the source may
    be overwritten as soon as the call returns (mlt_pool buffers are recycled immediately)."""
    rng = np.random.default_rng(nbytes)
    src = rng.integers(0, 256, nbytes, dtype=np.uint8)
    keep = src.copy()
    d = ctx.alloc(nbytes)
    abi.check(ctx.lib.mltcuda_upload_staged(ctx.h, d.ptr, src.ctypes.data_as(C.c_void_p), nbytes))
    src[:] = 0  # released / reused by the caller
    assert np.array_equal(ctx.get(d, (nbytes,)), keep)


def test_host_pool_recycles_exact_sizes(ctx):
    lib = ctx.lib
    a = lib.mltcuda_host_pool_alloc(1 << 20)
    b = lib.mltcuda_host_pool_alloc(1 << 20)
    assert a and b and a != b
    lib.mltcuda_host_pool_release(a)
    c = lib.mltcuda_host_pool_alloc(1 << 20)  # same size: the released buffer comes back
    assert c == a
    d = lib.mltcuda_host_pool_alloc(2 << 20)  # other size: a fresh one
    assert d not in (a, b)
    # a pooled buffer is pinned: a download into it is a plain async DMA
    dev = ctx.put(np.arange(256, dtype=np.uint8))
    abi.check(lib.mltcuda_download(ctx.h, c, dev.ptr, 256))
    ctx.sync()
    got = np.ctypeslib.as_array(C.cast(c, C.POINTER(C.c_uint8)), shape=(256,))
    assert np.array_equal(got, np.arange(256, dtype=np.uint8))
    for p in (b, c, d):
        lib.mltcuda_host_pool_release(p)


def test_wait_event_orders_two_contexts(ctx):
    """Work on a second context queued behind an event of the first sees the first one's result."""
    lib = ctx.lib
    other = Context(lib.mltcuda_ctx_device(ctx.h))
    n = 8 << 20
    a, b = ctx.alloc(n), ctx.alloc(n)
    ev = lib.mltcuda_event_create()
    for value in (0x11, 0x77):
        abi.check(lib.mltcuda_memset(ctx.h, a.ptr, value, n))
        abi.check(lib.mltcuda_event_record(ctx.h, ev))
        abi.check(lib.mltcuda_ctx_wait_event(other.h, ev))
        abi.check(lib.mltcuda_copy(other.h, b.ptr, a.ptr, n))
        other.sync()
        assert int(ctx.get(b, (n,)).min()) == value and int(ctx.get(b, (n,)).max()) == value
    lib.mltcuda_event_destroy(ev)


@pytest.mark.parametrize("nbytes", [1, 15, 16, 17, 1920 * 1080, 3840 * 2160 + 5])
def test_plane_opaque_query(ctx, nbytes):
    """is_opaque() (transition_luma.c:34-41) as a stream-ordered device query: all 0xff <=> opaque, any position."""
    lib = ctx.lib
    plane = np.full(nbytes, 0xff, np.uint8)
    d = ctx.put(plane)
    abi.check(lib.mltcuda_plane_opaque_query(ctx.h, d.ptr, nbytes))
    ctx.sync()
    assert lib.mltcuda_plane_opaque_result(ctx.h) == 1
    for pos in sorted({0, nbytes // 2, nbytes - 1}):
        p2 = plane.copy()
        p2[pos] = 0xfe
        d2 = ctx.put(p2)
        abi.check(lib.mltcuda_plane_opaque_query(ctx.h, d2.ptr, nbytes))
        ctx.sync()
        assert lib.mltcuda_plane_opaque_result(ctx.h) == 0, pos


def test_plane_opaque_query_into_keeps_verdicts_apart(ctx):
    """Two queries queued on ONE context before a single sync -- what two threads downloading images of the same
    context do -- each with a pinned word of its own: neither sees the other's verdict, in either order."""
    lib = ctx.lib
    n = 1920 * 1080
    opaque = ctx.put(np.full(n, 0xff, np.uint8))
    holed = np.full(n, 0xff, np.uint8)
    holed[n - 3] = 0x7f
    holed = ctx.put(holed)
    slots = [lib.mltcuda_host_pool_alloc(4) for _ in range(4)]
    words = [C.cast(s, C.POINTER(C.c_int)) for s in slots]
    try:
        for w in words:
            w[0] = 77  # stale contents must not matter
        for s, plane in zip(slots, (opaque, holed, holed, opaque)):
            abi.check(lib.mltcuda_plane_opaque_query_into(ctx.h, plane.ptr, n, s))
        ctx.sync()
        assert [w[0] for w in words] == [0, 1, 1, 0]
        # ragged sizes and the tail bytes, as in test_plane_opaque_query
        for nbytes in (1, 15, 17, 3840 * 2160 + 5):
            plane = np.full(nbytes, 0xff, np.uint8)
            abi.check(lib.mltcuda_plane_opaque_query_into(ctx.h, ctx.put(plane).ptr, nbytes, slots[0]))
            plane[nbytes - 1] = 0
            abi.check(lib.mltcuda_plane_opaque_query_into(ctx.h, ctx.put(plane).ptr, nbytes, slots[1]))
            ctx.sync()
            assert (words[0][0], words[1][0]) == (0, 1), nbytes
        pageable = (C.c_int * 1)()
        assert lib.mltcuda_plane_opaque_query_into(ctx.h, opaque.ptr, n, C.addressof(pageable)) != 0
        assert b"pinned" in lib.mltcuda_last_error()
    finally:
        for s in slots:
            lib.mltcuda_host_pool_release(s)


def test_graph_capture_replays_launches_on_two_contexts():
    """mltcuda_graph_*: a sequence recorded on one context, with a second context forked in and joined back through
    events, replays with one call and produces what the plain launches produce."""
    lib = abi.load()
    a, b = Context(0), Context(0)
    try:
        w, h = 256, 64
        d1, d2, ref1, ref2 = a.alloc(w * h * 2), b.alloc(w * h * 2), a.alloc(w * h * 2), a.alloc(w * h * 2)
        abi.check(lib.mltcuda_noise(a.h, ref1.ptr, w, h, 7, 1, 0))  # also warms the generator's tables
        abi.check(lib.mltcuda_noise(a.h, ref2.ptr, w, h, 8, 1, 0))
        abi.check(lib.mltcuda_noise(b.h, d2.ptr, w, h, 1, 1, 0))
        a.sync(), b.sync()
        fork, join = lib.mltcuda_event_create(), lib.mltcuda_event_create()
        abi.check(lib.mltcuda_graph_begin(a.h))
        lib.mltcuda_event_record(a.h, fork)
        abi.check(lib.mltcuda_ctx_wait_event(b.h, fork))
        abi.check(lib.mltcuda_noise(a.h, d1.ptr, w, h, 7, 1, 0))
        abi.check(lib.mltcuda_noise(b.h, d2.ptr, w, h, 8, 1, 0))
        lib.mltcuda_event_record(b.h, join)
        abi.check(lib.mltcuda_ctx_wait_event(a.h, join))
        g = lib.mltcuda_graph_end(a.h)
        assert g, lib.mltcuda_last_error()
        assert lib.mltcuda_graph_kernel_nodes(g) == 2
        n0 = a.launches()
        abi.check(lib.mltcuda_memset(a.h, d1.ptr, 0, w * h * 2))
        abi.check(lib.mltcuda_graph_launch(a.h, g))
        a.sync()
        assert a.launches() == n0 + 2
        assert np.array_equal(a.get(d1, (w * h * 2,)), a.get(ref1, (w * h * 2,)))
        assert np.array_equal(a.get(d2, (w * h * 2,)), a.get(ref2, (w * h * 2,)))
        lib.mltcuda_graph_destroy(g)
    finally:
        a.close(), b.close()
