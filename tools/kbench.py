#!/usr/bin/env python
"""tools/kbench.py -- per-kernel HBM roofline of every dedicated kernel (K1-K4, K6) at 2160p.

Each case launches one batched kernel over BATCH distinct frames (working set >> 126 MB L2),
timed with CUDA events on the launching stream after warm-up.  GB/s = ALGORITHMIC bytes
(SURVEY 8d per-pixel figures x pixels) / mean launch time; frac = vs MEASURED_PEAKS.json hbm_gbs.
Prints one JSON line per case; `--out FILE` also writes them to a file (profiles/).
Numbers printed under ncu are not bench values."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from mlt_b200 import Context, abi  # noqa: E402

W, H = 3840, 2160
NPX = W * H


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except (OSError, KeyError):
        return 6650.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--only", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    B = args.batch
    lib = abi.load()
    ctx = Context(0)
    pk = peak()
    yb, rb, ab_ = NPX * 2, NPX * 4, NPX
    # buffers: filled with noise so arithmetic sees varied bytes
    d_y0, d_y1 = ctx.alloc(B * yb), ctx.alloc(B * yb)
    d_r0, d_r1 = ctx.alloc(B * rb), ctx.alloc(B * rb)
    d_a0, d_a1 = ctx.alloc(B * ab_), ctx.alloc(B * ab_)
    # luma map: a horizontal ramp 0..65535 like demo/luma1.pgm (a wipe); with pos 0.4 / softness 0.2
    # about a fifth of every row lies inside the soft band (smoothstep polynomial), the rest is 0 / 1
    import numpy as np
    ramp = np.tile((np.arange(W, dtype=np.uint64) * 65535 // (W - 1)).astype(np.uint16), H)
    d_map = ctx.put(ramp.view(np.uint8))
    abi.check(lib.mltcuda_noise(ctx.h, d_y0.ptr, W, H, 1, B, yb))
    abi.check(lib.mltcuda_noise(ctx.h, d_y1.ptr, W, H, 1001, B, yb))
    abi.check(lib.mltcuda_noise(ctx.h, d_r0.ptr, W * 2, H, 2001, B, rb))
    abi.check(lib.mltcuda_noise(ctx.h, d_r1.ptr, W * 2, H, 3001, B, rb))
    abi.check(lib.mltcuda_noise(ctx.h, d_a0.ptr, W // 2, H, 4001, B, ab_))
    abi.check(lib.mltcuda_noise(ctx.h, d_a1.ptr, W // 2, H, 5001, B, ab_))
    ctx.sync()
    ev0, ev1 = lib.mltcuda_event_create(), lib.mltcuda_event_create()
    lines = []

    def run(name, fn, bytes_per_launch, note=""):
        if args.only and not any(o in name for o in args.only.split("|")):
            return
        for _ in range(3):
            fn()
        ctx.sync()
        lib.mltcuda_event_record(ctx.h, ev0)
        for _ in range(args.reps):
            fn()
        lib.mltcuda_event_record(ctx.h, ev1)
        ctx.sync()
        ms = lib.mltcuda_event_elapsed_ms(ev0, ev1) / args.reps
        gbs = bytes_per_launch / (ms * 1e-3) / 1e9
        line = {"kernel": name, "ms_per_launch": round(ms, 4), "frames_per_launch": B,
                "fps": round(B / (ms * 1e-3), 1), "algorithmic_bytes": int(bytes_per_launch),
                "achieved_gbs": round(gbs, 1), "peak_gbs": pk, "frac": round(gbs / pk, 4), "note": note}
        lines.append(line)
        print(json.dumps(line), flush=True)

    def call(name, desc):
        return lambda: abi.check(getattr(lib, name)(ctx.h, C.byref(desc)), name)

    # K1 imageconvert
    cv = abi.ConvertDesc(src_format=3, dst_format=2, width=W, height=H, src=d_y0.ptr, dst=d_r0.ptr, batch=B,
                         src_bstride=yb, dst_bstride=rb)
    run("K1 yuv422->rgba", call("mltcuda_convert", cv), B * NPX * 6)
    cv2 = abi.ConvertDesc(src_format=3, dst_format=2, width=W, height=H, src=d_y0.ptr, dst=d_r0.ptr, src_alpha=d_a0.ptr,
                          batch=B, src_bstride=yb, dst_bstride=rb, src_alpha_bstride=ab_)
    run("K1 yuv422+alpha->rgba", call("mltcuda_convert", cv2), B * NPX * 7)
    cv3 = abi.ConvertDesc(src_format=2, dst_format=3, width=W, height=H, src=d_r1.ptr, dst=d_y0.ptr, dst_alpha=d_a0.ptr,
                          batch=B, src_bstride=rb, dst_bstride=yb, dst_alpha_bstride=ab_)
    run("K1 rgba->yuv422+alpha", call("mltcuda_convert", cv3), B * NPX * 7)

    # the other cells of the matrix (generic pixel-pair kernel / 420p kernel); buffers reused as scratch
    bpp = {1: 3, 2: 4, 3: 2, 4: 1.5, 10: 8}
    for sf, df, nm in ((3, 4, "yuv422->yuv420p"), (4, 3, "yuv420p->yuv422"), (3, 1, "yuv422->rgb"), (1, 3, "rgb->yuv422"),
                       (1, 2, "rgb->rgba"), (2, 1, "rgba->rgb"), (4, 2, "yuv420p->rgba"), (4, 1, "yuv420p->rgb"),
                       (3, 10, "yuv422->rgba64"), (10, 3, "rgba64->yuv422"), (2, 10, "rgba->rgba64"), (10, 2, "rgba64->rgba")):
        w = W // 2 if 10 in (sf, df) else W  # rgba64 frames are W/2 wide in the rgba-sized buffers
        sbs, dbs = int(w * H * bpp[sf]), int(w * H * bpp[df])
        al = dict(dst_alpha=d_a0.ptr, dst_alpha_bstride=ab_) if sf in (2, 10) and df in (1, 3) else {}
        cvx = abi.ConvertDesc(**al, src_format=sf, dst_format=df, width=w, height=H, src=d_r1.ptr, dst=d_r0.ptr, batch=B,
                              src_bstride=sbs, dst_bstride=dbs)
        run(f"K1 {nm}", call("mltcuda_convert", cvx), B * (sbs + dbs))

    # K2 composite, full frame
    def comp(weight, ab=False, aa=False, luma=False, op=0):
        return abi.CompositeDesc(dest=d_y0.ptr, src=d_y1.ptr, alpha_b=d_a1.ptr if ab else None,
                                 alpha_a=d_a0.ptr if aa else None, luma=d_map.ptr if luma else None, width=W, rows=H,
                                 stride_dest=W * 2, stride_src=W * 2, stride_alpha_a=W, stride_alpha_b=W,
                                 weight=weight, softness=13107, step=30000, op=op, batch=B, dest_bstride=yb,
                                 src_bstride=yb, alpha_a_bstride=ab_, alpha_b_bstride=ab_, luma_bstride=0)
    def comp_rect(weight, x, y, w, h, sx=0, ab=False, aa=False, step=1):
        # B (its own w x h image, tightly packed from d_y1) over the rectangle (x, y, w, h) of A
        rows = h // step
        return abi.CompositeDesc(dest=d_y0.ptr + (y * W + x) * 2, src=d_y1.ptr + sx * 2,
                                 alpha_b=(d_a1.ptr + sx) if ab else None, alpha_a=(d_a0.ptr + y * W + x) if aa else None,
                                 luma=None, width=w, rows=rows, stride_dest=W * 2 * step, stride_src=w * 2 * step,
                                 stride_alpha_a=W * step, stride_alpha_b=w * step, weight=weight, softness=0, step=0, op=0,
                                 batch=B, dest_bstride=yb, src_bstride=yb, alpha_a_bstride=ab_, alpha_b_bstride=ab_, luma_bstride=0)
    rw, rh = 1537, 1081
    run("K2 composite 50% sub-rect at odd x (misaligned rows)", call("mltcuda_composite", comp_rect(32768, 101, 77, rw, rh)),
        B * rw * rh * 6)
    run("K2 composite 70% sub-rect odd x +alpha_a+alpha_b", call("mltcuda_composite", comp_rect(45875, 101, 77, rw, rh, ab=True, aa=True)),
        B * rw * rh * 9)
    run("K2 composite 50% sub-rect aligned (x=96, w=1536)", call("mltcuda_composite", comp_rect(32768, 96, 77, 1536, rh)),
        B * 1536 * rh * 6)
    run("K2 composite 70% sub-rect aligned +alpha_a+alpha_b", call("mltcuda_composite", comp_rect(45875, 96, 77, 1536, rh, ab=True, aa=True)),
        B * 1536 * rh * 9)
    run("K2 composite 50% one field (row step 2)", call("mltcuda_composite", comp_rect(32768, 0, 0, W, H, step=2)),
        B * NPX // 2 * 6)
    run("K2 composite 50% (sse2 case1)", call("mltcuda_composite", comp(32768)), B * NPX * 6)
    run("K2 composite 70% +alpha_b (case3)", call("mltcuda_composite", comp(45875, ab=True)), B * NPX * 7)
    run("K2 composite 70% +alpha_a+alpha_b (case7)", call("mltcuda_composite", comp(45875, ab=True, aa=True)), B * NPX * 9)
    run("K2 composite blit (case0)", call("mltcuda_composite", comp(65536)), B * NPX * 4)
    run("K2 composite luma wipe (scalar)", call("mltcuda_composite", comp(32768, luma=True)), B * NPX * 6 + NPX * 2,
        "luma map shared by the batch")
    run("K2 composite xor +alpha (scalar)", call("mltcuda_composite", comp(32768, ab=True, aa=True, op=3)), B * NPX * 9)

    # K3 luma
    def luma(fmt, mapped=True, translucent=0, planes=False):
        img0, img1, bs = (d_y0, d_y1, yb) if fmt == 3 else (d_r0, d_r1, rb)
        w = W if fmt != 10 else W // 2
        extra = dict(alpha_dest=d_a0.ptr, alpha_src=d_a1.ptr, alpha_dest_bstride=ab_, alpha_src_bstride=ab_) if planes else {}
        return abi.LumaDesc(**extra, format=fmt, dest=img0.ptr, src=img1.ptr, map=d_map.ptr if mapped else None, luma_width=W,
                            luma_height=H, width=w, height=H, width_dest=w, height_dest=H, width_src=w, height_src=H,
                            pos=0.4, frame_delta=0.004, softness=0.2, field_order=-1, invert=0, alpha_over=1,
                            translucent=translucent, batch=B, dest_bstride=bs, src_bstride=bs)
    run("K3 luma wipe yuv422 opaque", call("mltcuda_luma", luma(3)), B * NPX * 6 + NPX * 2)
    run("K3 luma wipe yuv422 translucent(float, alpha planes)", call("mltcuda_luma", luma(3, translucent=1, planes=True)),
        B * NPX * 9 + NPX * 2)
    run("K3 dissolve yuv422 translucent(float, alpha planes)", call("mltcuda_luma", luma(3, mapped=False, translucent=1, planes=True)),
        B * NPX * 9)
    run("K3 dissolve yuv422 opaque (cross-fade)", call("mltcuda_luma", luma(3, mapped=False)), B * NPX * 6)
    run("K3 dissolve yuv422 opaque alpha planes", call("mltcuda_luma", luma(3, mapped=False, planes=True)), B * NPX * 9)
    run("K3 dissolve rgba opaque", call("mltcuda_luma", luma(2, mapped=False)), B * NPX * 12)
    run("K3 luma wipe rgba opaque", call("mltcuda_luma", luma(2)), B * NPX * 12 + NPX * 2)
    run("K3 luma wipe rgba64 opaque", call("mltcuda_luma", luma(10)), B * (NPX // 2) * 24 + NPX)
    run("K3 luma wipe rgba64 translucent(float)", call("mltcuda_luma", luma(10, translucent=1)), B * (NPX // 2) * 24 + NPX)
    run("K3 dissolve rgba64 translucent(float)", call("mltcuda_luma", luma(10, mapped=False, translucent=1)), B * (NPX // 2) * 24)
    lf = luma(3)
    lf.field_order = 0
    run("K3 luma wipe yuv422 opaque, interlaced (field_order=0)", call("mltcuda_luma", lf), B * NPX * 6 + NPX * 2)
    run("K3 luma wipe rgba translucent(float)", call("mltcuda_luma", luma(2, translucent=1)), B * NPX * 12 + NPX * 2)
    run("K3 dissolve rgba translucent(float)", call("mltcuda_luma", luma(2, mapped=False, translucent=1)), B * NPX * 12)

    # K4 affine: B covers all of A (scale 1, 3 degree rotation about the centre, B same size)
    import math

    def aff(fmt, interp, deg=3.0, scale=1.0):
        w = W if fmt == 2 else W // 2
        c, s = math.cos(math.radians(deg)), math.sin(math.radians(deg))
        d = abi.AffineDesc(format=fmt, a_image=d_r0.ptr, b_image=d_r1.ptr, a_width=w, a_height=H, b_width=w, b_height=H,
                           lower_x=-w / 2.0, lower_y=-H / 2.0, dz=1.0, mix=0.8, x_offset=w / 2.0, y_offset=H / 2.0,
                           minima=0.0, xmax=w - 1.0, ymax=H - 1.0, b_alpha=0, interp=interp, batch=B, a_bstride=rb,
                           b_bstride=rb)
        m = [[c * scale, -s * scale, 0.0], [s * scale, c * scale, 0.0], [0.0, 0.0, 1.0]]
        for i in range(3):
            for j in range(3):
                d.matrix[i][j] = m[i][j]
        return d

    def covered(d, w):
        import numpy as np
        xs = d.lower_x + np.arange(0, w, 4, dtype=np.float64)
        ys = d.lower_y + np.arange(0, H, 4, dtype=np.float64)
        X, Y = np.meshgrid(xs, ys)
        dx = (d.matrix[0][0] * X + d.matrix[0][1] * Y + d.matrix[0][2]) / d.dz + d.x_offset
        dy = (d.matrix[1][0] * X + d.matrix[1][1] * Y + d.matrix[1][2]) / d.dz + d.y_offset
        return int(((dx >= d.minima) & (dx <= d.xmax) & (dy >= d.minima) & (dy <= d.ymax)).sum()) * 16

    for nm, fmt, it in (("rgba bilinear", 2, 1), ("rgba nearest", 2, 0), ("rgba bicubic", 2, 2),
                        ("rgba64 bilinear", 10, 1), ("rgba64 nearest", 10, 0), ("rgba64 bicubic", 10, 2)):
        d = aff(fmt, it)
        w = W if fmt == 2 else W // 2
        cov = covered(d, w)
        run(f"K4 affine {nm} rot3", call("mltcuda_affine", d), B * cov * (12 if fmt == 2 else 24),
            f"covered {cov / (w * H):.3f} of dest")

    # K6 producers
    run("K6 noise yuv422", lambda: abi.check(lib.mltcuda_noise(ctx.h, d_y1.ptr, W, H, 7, B, yb)), B * NPX * 2)
    hb = (W // 2) * (H // 2) * 2
    d_half = ctx.alloc(B * hb)
    run("K6 rescale_nn yuv422 2160p->1080p",
        lambda: abi.check(lib.mltcuda_rescale_yuv422(ctx.h, d_half.ptr, W // 2, H // 2, d_y0.ptr, W, H, B, hb, yb)),
        B * (W // 2) * (H // 2) * 4, "2 B/px written + 2 B/px sampled (every other pixel of every other row)")
    run("K6 mirror flip yuv422 (16 frames, one launch each)",
        lambda: [abi.check(lib.mltcuda_mirror_flip(ctx.h, d_y1.ptr + i * yb, None, d_y0.ptr + i * yb, None, W, H)) for i in range(B)],
        B * NPX * 4)
    if args.out:
        with open(args.out, "w") as f:
            for ln in lines:
                f.write(json.dumps(ln) + "\n")
    ctx.close()


if __name__ == "__main__":
    main()
