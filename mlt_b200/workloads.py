"""Host-side parameter derivation for the benchmark configurations of BASELINE.json (cfg1..cfg4).

This restates, for the concrete melt command lines of SURVEY 8(d), what the reference's host code
computes per frame before it touches pixels -- composite_calculate / get_b_frame_image
(src/modules/core/transition_composite.c:730-941) and the affine transition's get_image
(src/modules/plus/transition_affine.c:414-729, get_affine :268-326, affine_* :93-210) -- as plain
Python floats (= C doubles, same operation order).  The MLT module (mlt_b200/module) does the same
through mlt_properties inside MLT; this file exists so bench.py and the tests can drive the C-ABI
with exactly the per-frame parameters the reference would use (tests/test_cfg_vs_reference.py
checks the derivation against frames rendered by the reference itself).
"""
import ctypes as C
import math

from . import abi

W_UHD, H_UHD = 3840, 2160
RGBA, YUV422 = 2, 3


def rint(x):
    return int(round(x))  # python round() is round-half-even like rint()


def lerp_rect(k0, r0, k1, r1, pos):
    """mlt_property_anim_get_rect, linear keyframes (src/framework/mlt_property.c:1265,1385-1389)."""
    if pos <= k0:
        return list(r0)
    if pos >= k1:
        return list(r1)
    t = (pos - k0) / float(k1 - k0)
    return [a + (b - a) * t for a, b in zip(r0, r1)]


def composite_geometry(rect_pct, W, H, b_real_w, b_real_h):
    """composite_calculate (:866-941) + get_b_frame_image default branch (:760-808) for
    aligned=1, distort=0, fill=0, sar 1:1, geometry given in % with opacity fraction.
    Returns (geometry dict for mltcuda_composite_setup, B width, B height)."""
    x, y, w, h, o = rect_pct
    item = dict(x=x * W, y=y * H, w=w * W, h=h * H, o=100.0 * min(o, 1.0))
    nw, nh = W, H
    norm_w, norm_h = int(item["w"]), int(item["h"])
    sw, sh = b_real_w, b_real_h
    if sw > norm_w:
        sh = (sh * norm_w) // sw  # int arithmetic, :775
        sw = norm_w
    if sh > norm_h:
        sw = (sw * norm_h) // sh
        sh = norm_h
    # alignment_calculate with halign = valign = 0 adds nothing
    bw = rint(sw * W / nw)
    bw -= bw % 2
    bh = rint(sh * H / nh)
    g = dict(item, nw=nw, nh=nh, sw=bw, sh=sh, x_src=0, y_src=0)  # :826-827 sw = *width
    return g, bw, bh


def mat_identity():
    return [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]


def mat_multiply(a, m):
    """affine_multiply (:108-127): a <- a x m^T"""
    return [[a[i][0] * m[j][0] + a[i][1] * m[j][1] + a[i][2] * m[j][2] for j in range(3)] for i in range(3)]


def affine_params(W, H, rect_px, position, rotate_x_per_frame, b_w, b_h, interp="bilinear"):
    """Per-frame sliced_desc of the affine transition (:616-714) for fill=1, distort=0, sar 1:1,
    B requested at its native size, keyed=0, only rotate_x animated, no offsets/shear/alignment."""
    x, y, w, h, o = rect_px
    mix = 1.0 if o is None else min(o, 1.0)
    geom_scale_x, geom_scale_y = b_w / w, b_h / h
    d = dict(lower_x=-(x + w / 2.0), lower_y=-(y + h / 2.0), mix=mix, x_offset=b_w / 2.0, y_offset=b_h / 2.0,
             b_alpha=0, minima=0.0, xmax=b_w - 1.0, ymax=b_h - 1.0, b_width=b_w, b_height=b_h)
    m = mat_identity()
    ang = 0.0 + rotate_x_per_frame * position
    c, s = math.cos(ang * math.pi / 180), math.sin(ang * math.pi / 180)
    m = mat_multiply(m, [[c, 0 - s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])  # affine_rotate_x
    m = mat_multiply(m, [[1.0, 0.0, 0 - 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]])  # rotate_y(0)
    m = mat_multiply(m, [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, -0.0, 1.0]])  # rotate_z(0)
    m = mat_multiply(m, [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]])  # shear(0)
    d["dz"] = m[2][2]
    b_dar, consumer_dar = b_w / float(b_h), W / float(H)
    if b_dar > consumer_dar:
        sx, sy = geom_scale_x, geom_scale_x
    else:
        sx, sy = geom_scale_y, geom_scale_y
    m = mat_multiply(m, [[sx, 0.0, 0.0], [0.0, sy, 0.0], [0.0, 0.0, 1.0]])  # affine_scale
    d["matrix"] = m
    if interp in ("nearest", "neighbor", "tiles", "fast_bilinear", None):
        d["interp"], d["minima"], d["xmax"], d["ymax"] = 0, d["minima"] - 0.5, d["xmax"] + 0.49, d["ymax"] + 0.49
    elif interp == "bilinear":
        d["interp"] = 1
    else:
        d["interp"], d["minima"] = 2, d["minima"] - 1
    return d


# ---- the configurations ---------------------------------------------------------------------------

CFG4 = dict(
    name="cfg4_uhd_2160p_60_4track",
    comp1=(0.05, 0.05, 0.40, 0.40, 0.80),
    affine_k0=(0, (0.20, 0.20, 0.50, 0.50, 0.90)), affine_k1=(599, (0.30, 0.25, 0.60, 0.60, 0.70)),
    rotate_x=0.5, colour=(0x30, 0x60, 0xc0, 0xa0),
    comp2=(0.55, 0.55, 0.40, 0.40, 1.00),
)


def cfg4_frame(position, W=W_UHD, H=H_UHD, cfg=CFG4):
    """Everything the three transitions of cfg4 need at timeline position `position`."""
    g1, b1w, b1h = composite_geometry(cfg["comp1"], W, H, W, H)
    g2, b2w, b2h = composite_geometry(cfg["comp2"], W, H, W, H)
    (k0, r0), (k1, r1) = cfg["affine_k0"], cfg["affine_k1"]
    r = lerp_rect(k0, r0, k1, r1, position)
    rect_px = [r[0] * W, r[1] * H, r[2] * W, r[3] * H, r[4]]
    aff = affine_params(W, H, rect_px, position, cfg["rotate_x"], W, H, "bilinear")
    return dict(W=W, H=H, comp1=(g1, b1w, b1h), comp2=(g2, b2w, b2h), affine=aff)


def shard_frames(rank, world, frames_per_rank, step=0):
    """Frame-parallel sharding of a timeline over `world` GPUs: frames are independent,
    so rank r renders the contiguous block [r*F, (r+1)*F) of every step's world*F frames.  No data
    exchange is needed; order is restored by frame index."""
    base = step * world * frames_per_rank + rank * frames_per_rank
    return list(range(base, base + frames_per_rank))


# ---- C-ABI descriptors for those parameters (ctypes mirrors of include/mltcuda.h) ---------------------

def geometry_struct(g):
    return abi.CompositeGeometry(x=g["x"], y=g["y"], w=g["w"], h=g["h"], o=g["o"], nw=g["nw"], nh=g["nh"],
                                 sw=g["sw"], sh=g["sh"], x_src=g["x_src"], y_src=g["y_src"])


def affine_desc(k, a_ptr, b_ptr, W, H, batch=1, a_bs=0, b_bs=0):
    d = abi.AffineDesc(format=RGBA, a_image=a_ptr, b_image=b_ptr, a_width=W, a_height=H, b_width=k["b_width"],
                       b_height=k["b_height"], lower_x=k["lower_x"], lower_y=k["lower_y"], dz=k["dz"], mix=k["mix"],
                       x_offset=k["x_offset"], y_offset=k["y_offset"], minima=k["minima"], xmax=k["xmax"],
                       ymax=k["ymax"], b_alpha=k["b_alpha"], interp=k["interp"], batch=batch, a_bstride=a_bs,
                       b_bstride=b_bs)
    for i in range(3):
        for j in range(3):
            d.matrix[i][j] = float(k["matrix"][i][j])
    return d


def composite_desc(ctx, g, dest_ptr, W, H, src_ptr, bw, bh, alpha_a_ptr=None, batch=1, dest_bs=0, src_bs=0, aa_bs=0,
                   alpha_b_ptr=None, luma_ptr=None, softness=0.0):
    d = abi.CompositeDesc(batch=batch, dest_bstride=dest_bs, src_bstride=src_bs, alpha_a_bstride=aa_bs)
    gs = geometry_struct(g)
    skip = ctx.lib.mltcuda_composite_setup(C.byref(d), dest_ptr, W, H, src_ptr, bw, bh, alpha_b_ptr, alpha_a_ptr, C.byref(gs),
                                           -1, luma_ptr, softness)
    assert not skip
    return d


def tractor_desc(ctx, f, in_ptr, out_ptr, out_alpha_ptr, b1_ptr, bc_ptr, b3_ptr, batch=1, bs=None):
    """The cfg4 op list for mltcuda_tractor.  bs = dict of batch strides."""
    W, H = f["W"], f["H"]
    (g1, b1w, b1h), (g2, b2w, b2h), k = f["comp1"], f["comp2"], f["affine"]
    bs = bs or {}
    ops = (abi.TractorOp * 5)()
    ops[0].kind = abi.TOP_COMPOSITE
    ops[0].u.composite = composite_desc(ctx, g1, out_ptr, W, H, b1_ptr, b1w, b1h, src_bs=bs.get("b1", 0))
    ops[1].kind, ops[1].to_format = abi.TOP_CONVERT, RGBA
    ops[2].kind = abi.TOP_AFFINE
    ops[2].u.affine = affine_desc(k, None, bc_ptr, W, H, b_bs=bs.get("bc", 0))
    ops[3].kind, ops[3].to_format = abi.TOP_CONVERT, YUV422
    ops[4].kind = abi.TOP_COMPOSITE
    ops[4].u.composite = composite_desc(ctx, g2, out_ptr, W, H, b3_ptr, b2w, b2h, src_bs=bs.get("b3", 0))
    d = abi.TractorDesc(width=W, height=H, in_format=YUV422, out_format=YUV422, in_=in_ptr, out=out_ptr,
                        out_alpha=out_alpha_ptr, n_ops=5, ops=ops, batch=batch, in_bstride=bs.get("in", 0),
                        out_bstride=bs.get("out", 0), out_alpha_bstride=bs.get("out_alpha", 0))
    d._keep = ops
    return d


# ---- cfg1 .. cfg3 (1080p): the single-transition configurations of BASELINE.json ------------------------

CFG1B = dict(name="cfg1b_1080p_composite_blend", geometry=(0.0, 0.0, 1.0, 1.0, 0.5))
CFG3 = dict(name="cfg3_1080p_affine_rgba", k0=(0, (0.10, 0.10, 0.50, 0.50, 1.0)), k1=(249, (0.30, 0.30, 0.70, 0.70, 0.6)),
            rotate_x=1.0)


def cfg3_frame(position, W=1920, H=1080, cfg=CFG3):
    (k0, r0), (k1, r1) = cfg["k0"], cfg["k1"]
    r = lerp_rect(k0, r0, k1, r1, position)
    rect_px = [r[0] * W, r[1] * H, r[2] * W, r[3] * H, r[4]]
    return affine_params(W, H, rect_px, position, cfg["rotate_x"], W, H, "bilinear")


def interleaved_frames(rank, world, frames_per_rank, window_ranks=8):
    """Frame-parallel sharding with EQUAL work per GPU: the cost of a cfg4 frame drifts along the timeline (the
    keyframed affine rectangle grows), so contiguous blocks would give later ranks heavier frames and a max-over-ranks
    timing would read that as a scaling loss.  The benchmark therefore always works on the same window of
    frames_per_rank x max(world, window_ranks) timeline frames, dealt round-robin: rank r renders frames
    r, r + S, r + 2S, ... with S = max(world, window_ranks) -- every rank (and the 1-GPU run) sees the same cost profile."""
    stride = max(world, window_ranks)
    return [i * stride + rank for i in range(frames_per_rank)]
