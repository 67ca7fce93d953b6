"""cfg4 (4-track tractor: composite + affine + composite with the imageconvert round trip) built
three ways on the same synthetic inputs: CPU oracle chain, GPU one-launch-per-op, GPU fused."""
import ctypes as C

import numpy as np

import oracle
from mlt_b200 import abi, workloads

RGBA, YUV422 = 2, 3


def flip_yuv422(img, w, h):
    """core filter mirror:flip (src/modules/core/filter_mirror.c:199-218), even widths."""
    a = img.reshape(h, w // 2, 4)[:, ::-1, :][:, :, [2, 1, 0, 3]]
    return np.ascontiguousarray(a).reshape(-1)


def colour_b_rgba(w, h, rgba):
    """colour: producer asked for yuv422 (+ alpha plane), then imageconvert -> rgba, as the reference's
    normaliser chain does for the affine B track (melt -debug)."""
    yuv, al = oracle.colour(YUV422, w, h, *rgba)
    _, out, _ = oracle.convert(YUV422, RGBA, w, h, yuv, al)
    return out


def inputs(position, W, H, interp="bilinear"):
    f = workloads.cfg4_frame(position, W, H)
    if interp != "bilinear":  # the same frame with another sampler (consumer.rescale=nearest / bicubic)
        cfg = workloads.CFG4
        (k0, r0), (k1, r1) = cfg["affine_k0"], cfg["affine_k1"]
        r = workloads.lerp_rect(k0, r0, k1, r1, position)
        f["affine"] = workloads.affine_params(W, H, [r[0] * W, r[1] * H, r[2] * W, r[3] * H, r[4]], position, cfg["rotate_x"], W, H, interp)
    (g1, b1w, b1h), (g2, b2w, b2h) = f["comp1"], f["comp2"]
    return dict(f=f, a=oracle.noise(W, H, position), b1=flip_yuv422(oracle.noise(b1w, b1h, position), b1w, b1h),
                bc=colour_b_rgba(W, H, workloads.CFG4["colour"]), b3=oracle.noise(b2w, b2h, position))


def _item(g):
    return [g["x"], g["y"], g["w"], g["h"], g["o"]], [g["nw"], g["nh"], g["sw"], g["sh"], g["x_src"], g["y_src"]]


def oracle_chain(inp):
    f = inp["f"]
    W, H = f["W"], f["H"]
    (g1, b1w, b1h), (g2, b2w, b2h), k = f["comp1"], f["comp2"], f["affine"]
    a1, _, _ = oracle.composite_yuv(inp["a"], W, H, inp["b1"], b1w, b1h, None, None, *_item(g1))
    _, rgba, _ = oracle.convert(YUV422, RGBA, W, H, a1)
    rgba2, covered = oracle.affine(8, rgba, inp["bc"], W, H, k["b_width"], k["b_height"], k["matrix"], k["lower_x"],
                                   k["lower_y"], k["dz"], k["mix"], k["x_offset"], k["y_offset"], k["minima"],
                                   k["xmax"], k["ymax"], k["b_alpha"], k["interp"])
    _, yuv, alpha = oracle.convert(RGBA, YUV422, W, H, rgba2, want_alpha=True)
    out, out_a, _ = oracle.composite_yuv(yuv, W, H, inp["b3"], b2w, b2h, None, alpha, *_item(g2))
    return dict(out=out, alpha=out_a, stage1=a1, stage2=rgba2, covered=covered)


from mlt_b200.workloads import affine_desc, composite_desc, geometry_struct, tractor_desc  # noqa: E402,F401


def gpu_unfused(ctx, inp):
    """One launch per reference pass: composite, convert, affine, convert, composite."""
    f = inp["f"]
    W, H = f["W"], f["H"]
    (g1, b1w, b1h), (g2, b2w, b2h), k = f["comp1"], f["comp2"], f["affine"]
    d_a, d_b1, d_bc, d_b3 = ctx.put(inp["a"]), ctx.put(inp["b1"]), ctx.put(inp["bc"]), ctx.put(inp["b3"])
    d_rgba, d_alpha = ctx.alloc(W * H * 4), ctx.alloc(W * H)
    ctx.call("mltcuda_composite", composite_desc(ctx, g1, d_a.ptr, W, H, d_b1.ptr, b1w, b1h))
    ctx.call("mltcuda_convert", abi.ConvertDesc(src_format=YUV422, dst_format=RGBA, width=W, height=H, src=d_a.ptr,
                                                dst=d_rgba.ptr, batch=1))
    ctx.call("mltcuda_affine", affine_desc(k, d_rgba.ptr, d_bc.ptr, W, H))
    ctx.call("mltcuda_convert", abi.ConvertDesc(src_format=RGBA, dst_format=YUV422, width=W, height=H, src=d_rgba.ptr,
                                                dst=d_a.ptr, dst_alpha=d_alpha.ptr, batch=1))
    ctx.call("mltcuda_composite", composite_desc(ctx, g2, d_a.ptr, W, H, d_b3.ptr, b2w, b2h, d_alpha.ptr))
    return dict(out=ctx.get(d_a, (W * H * 2,)), alpha=ctx.get(d_alpha, (W * H,)))


def gpu_fused(ctx, inp):
    f = inp["f"]
    W, H = f["W"], f["H"]
    d_a, d_b1, d_bc, d_b3 = ctx.put(inp["a"]), ctx.put(inp["b1"]), ctx.put(inp["bc"]), ctx.put(inp["b3"])
    d_out, d_alpha = ctx.alloc(W * H * 2), ctx.alloc(W * H)
    ctx.call("mltcuda_tractor", tractor_desc(ctx, f, d_a.ptr, d_out.ptr, d_alpha.ptr, d_b1.ptr, d_bc.ptr, d_b3.ptr))
    return dict(out=ctx.get(d_out, (W * H * 2,)), alpha=ctx.get(d_alpha, (W * H,)))
