"""Parity against the committed golden fixtures (tests/golden/, generated from the compiled
reference by tests/golden/make_golden.py).  These need neither /root/reference nor oracle/_ref.
CPU tier: the oracle reproduces every fixture.  GPU tier: the CUDA path (through the C-ABI)
reproduces them too -- bit-exact, including the float samplers."""
import json
import os
import sys

import numpy as np
import pytest

import oracle
import pipeline

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden as mg  # noqa: E402

G = os.path.join(HERE, "golden")
LINES = np.load(os.path.join(G, "composite_line_yuv.npz"))
INTERP = np.load(os.path.join(G, "interp.npz"))
FRAMES = json.load(open(os.path.join(G, "frames.json")))["frames"]


def fnv1a64(buf):
    """FNV-1a 64 as oracle/mlt_dump.c computes it (vectorised per 4 KiB is not possible: sequential)."""
    import ctypes as C
    lib = oracle.lib()
    if hasattr(lib, "orc_fnv1a64"):
        lib.orc_fnv1a64.restype = C.c_uint64
        return "%016x" % lib.orc_fnv1a64(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size))
    h = 1469598103934665603
    for b in buf.tobytes():
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


def test_oracle_reproduces_golden_composite_lines():
    rng = np.random.default_rng(mg.SEED)
    for name, w, weight, ab, aa, lum, soft, step in mg.line_cases():
        dest, src, alpha_b, alpha_a, luma = mg.line_inputs(rng, w, ab, aa, lum)
        d, a = oracle.composite_line(dest, src, w, alpha_b, alpha_a, weight, luma, soft, step)
        assert np.array_equal(d, LINES[name + "/dest"]), name
        if aa:
            assert np.array_equal(a, LINES[name + "/alpha_a"]), name


def test_oracle_reproduces_golden_interp_bit_exact():
    rng = np.random.default_rng(mg.SEED + 1)
    for bits in (8, 16):
        for kind in (0, 1, 2):
            exp = INTERP[f"b{bits}_k{kind}"]
            for i in range(exp.shape[0]):
                s, d, w, h, x, y, o, atop = mg.interp_inputs(rng, bits)
                got = oracle.interp_px(bits, kind, s, w, h, x, y, o, d, atop)
                assert np.array_equal(got, exp[i]), (bits, kind, i)


def test_oracle_chain_reproduces_golden_frame_hashes():
    W, H = 1920, 1080
    a = oracle.noise(W, H, 10)
    b = pipeline.flip_yuv422(oracle.noise(W, H, 10), W, H)
    exp, _, _ = oracle.composite_yuv(a, W, H, b, W, H, None, None, [0, 0, W, H, 50.0], [W, H, W, H, 0, 0])
    assert fnv1a64(exp) == FRAMES["cfg1b_blend"]["image_fnv1a64"] == "13c489d896346a6e"
    for frame, key in ((0, "cfg4_1080p_f0"), (37, "cfg4_1080p_f37")):
        res = pipeline.oracle_chain(pipeline.inputs(frame, W, H))
        assert fnv1a64(res["out"]) == FRAMES[key]["image_fnv1a64"], key
        assert fnv1a64(res["alpha"]) == FRAMES[key]["alpha_fnv1a64"], key


@pytest.mark.gpu
def test_gpu_reproduces_golden_composite_lines(ctx):
    from mlt_b200 import abi
    rng = np.random.default_rng(mg.SEED)
    for name, w, weight, ab, aa, lum, soft, step in mg.line_cases():
        dest, src, alpha_b, alpha_a, luma = mg.line_inputs(rng, w, ab, aa, lum)
        d_d, d_s = ctx.put(dest), ctx.put(src)
        d_ab = ctx.put(alpha_b) if ab else None
        d_aa = ctx.put(alpha_a) if aa else None
        d_l = ctx.put(luma) if lum else None
        desc = abi.CompositeDesc(dest=d_d.ptr, src=d_s.ptr, alpha_b=d_ab.ptr if ab else None,
                                 alpha_a=d_aa.ptr if aa else None, luma=d_l.ptr if lum else None, width=w, rows=1,
                                 stride_dest=2 * w, stride_src=2 * w, stride_alpha_a=w, stride_alpha_b=w, weight=weight,
                                 softness=soft, step=step, batch=1)
        ctx.call("mltcuda_composite", desc)
        assert np.array_equal(ctx.get(d_d, dest.shape), LINES[name + "/dest"]), name
        if aa:
            assert np.array_equal(ctx.get(d_aa, alpha_a.shape), LINES[name + "/alpha_a"]), name


@pytest.mark.gpu
def test_gpu_fused_tractor_reproduces_golden_cfg4_hashes(ctx):
    W, H = 1920, 1080
    for frame, key in ((0, "cfg4_1080p_f0"), (37, "cfg4_1080p_f37")):
        fus = pipeline.gpu_fused(ctx, pipeline.inputs(frame, W, H))
        assert fnv1a64(fus["out"]) == FRAMES[key]["image_fnv1a64"], key
        assert fnv1a64(fus["alpha"]) == FRAMES[key]["alpha_fnv1a64"], key


@pytest.mark.gpu
def test_gpu_fused_tractor_reproduces_golden_cfg4_hashes_at_2160p(ctx):
    """The full BASELINE size: frames the compiled reference rendered on uhd_2160p_60 (tests/golden/frames.json)."""
    W, H = 3840, 2160
    for frame, key in ((300, "cfg4_2160p_f300"), (599, "cfg4_2160p_f599")):
        assert FRAMES[key]["width"] == W and FRAMES[key]["height"] == H
        fus = pipeline.gpu_fused(ctx, pipeline.inputs(frame, W, H))
        assert fnv1a64(fus["out"]) == FRAMES[key]["image_fnv1a64"], key
        assert fnv1a64(fus["alpha"]) == FRAMES[key]["alpha_fnv1a64"], key
