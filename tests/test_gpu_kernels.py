"""GPU parity: every CUDA kernel, called through the C-ABI (include/mltcuda.h), against the CPU
oracle (oracle/mlt_oracle.c) on the same seeded inputs.  Integer paths must be bit-exact; float
paths (affine samplers, translucent luma/dissolve) within +-1 LSB."""
import ctypes as C

import numpy as np
import pytest

import oracle
from mlt_b200 import abi

pytestmark = pytest.mark.gpu

RGB, RGBA, YUV422, YUV420P, RGBA64 = 1, 2, 3, 4, 10


def rnd(rng, n, dtype=np.uint8):
    hi = 256 if dtype == np.uint8 else 65536
    return rng.integers(0, hi, n).astype(dtype)


# ------------------------------------------------------------------------------------------------
# K1 imageconvert

CELLS = [(YUV422, RGBA), (YUV422, RGB), (RGBA, YUV422), (RGB, YUV422), (YUV420P, YUV422), (YUV420P, RGB),
         (YUV420P, RGBA), (YUV422, YUV420P), (RGB, RGBA), (RGBA, RGB), (RGBA, RGBA64), (RGBA64, RGBA),
         (YUV422, RGBA64), (RGBA64, YUV422)]


@pytest.mark.parametrize("sf,df", CELLS)
@pytest.mark.parametrize("w,h", [(1920, 16), (64, 6), (70, 5), (33, 4), (2, 2), (8, 1)])
@pytest.mark.parametrize("with_alpha", [False, True])
def test_convert(ctx, sf, df, w, h, with_alpha):
    if YUV420P in (sf, df) and (w % 2 or h % 2):
        pytest.skip("yuv420p needs even sizes")
    rng = np.random.default_rng(w * 131 + h * 7 + sf * 3 + df)
    src = rnd(rng, oracle.image_bytes(sf, w, h))
    if sf == YUV420P:
        # the reference addresses chroma rows with stride*line/2 (not stride*(line/2)), so odd lines
        # read up to half a chroma row past the V plane: give both sides the same defined padding
        src = np.concatenate([src, np.zeros(w, np.uint8)])
    sa = rnd(rng, w * h) if with_alpha and df in (RGBA, RGBA64) and sf in (YUV422, YUV420P, RGB) else None
    want_alpha = sf in (RGBA, RGBA64) and df in (YUV422, RGB)
    init = rnd(rng, oracle.image_bytes(df, w, h))  # bytes the reference leaves untouched must survive
    rc, exp, exp_a = oracle.convert(sf, df, w, h, src, sa, want_alpha, dst_init=init)
    assert rc == 0
    d_src, d_dst = ctx.put(src), ctx.put(init)
    d_sa = ctx.put(sa) if sa is not None else None
    d_da = ctx.put(np.zeros(w * h, np.uint8)) if want_alpha else None
    desc = abi.ConvertDesc(src_format=sf, dst_format=df, width=w, height=h, src=d_src.ptr, dst=d_dst.ptr,
                           src_alpha=d_sa.ptr if d_sa else None, dst_alpha=d_da.ptr if d_da else None, batch=1)
    ctx.call("mltcuda_convert", desc)
    got = ctx.get(d_dst, exp.shape)
    assert np.array_equal(got, exp), f"{np.flatnonzero(got != exp)[:8]}"
    if want_alpha:
        assert np.array_equal(ctx.get(d_da, exp_a.shape), exp_a)


def test_convert_null_cells_rejected(ctx):
    d = ctx.put(np.zeros(1024, np.uint8))
    for sf, df in [(RGB, YUV420P), (RGBA, YUV420P), (RGB, RGBA64), (YUV422, YUV422)]:
        desc = abi.ConvertDesc(src_format=sf, dst_format=df, width=4, height=4, src=d.ptr, dst=d.ptr, dst_alpha=d.ptr)
        assert ctx.lib.mltcuda_convert(ctx.h, C.byref(desc)) == abi.E_UNSUPPORTED


def test_convert_batched_strided(ctx):
    rng = np.random.default_rng(5)
    w, h, n = 256, 8, 3
    src = rnd(rng, n * w * h * 2)
    d_src, d_dst = ctx.put(src), ctx.put(np.zeros(n * w * h * 4, np.uint8))
    desc = abi.ConvertDesc(src_format=YUV422, dst_format=RGBA, width=w, height=h, src=d_src.ptr, dst=d_dst.ptr,
                           batch=n, src_bstride=w * h * 2, dst_bstride=w * h * 4)
    ctx.call("mltcuda_convert", desc)
    got = ctx.get(d_dst, (n, w * h * 4))
    for i in range(n):
        _, exp, _ = oracle.convert(YUV422, RGBA, w, h, src[i * w * h * 2:(i + 1) * w * h * 2])
        assert np.array_equal(got[i], exp)


# ------------------------------------------------------------------------------------------------
# K2 composite

def run_composite_rows(ctx, dest, src, ab, aa, luma, off, width, rows, sd, ss, saa, sab, weight, soft, step, op=0,
                       scalar_only=0):
    d_dest, d_src = ctx.put(dest), ctx.put(src)
    d_ab = ctx.put(ab) if ab is not None else None
    d_aa = ctx.put(aa) if aa is not None else None
    d_l = ctx.put(luma) if luma is not None else None
    desc = abi.CompositeDesc(dest=d_dest.at(off["dest"]), src=d_src.at(off["src"]),
                             alpha_b=d_ab.at(off.get("alpha_b", 0)) if d_ab else None,
                             alpha_a=d_aa.at(off.get("alpha_a", 0)) if d_aa else None,
                             luma=d_l.at(2 * off.get("luma", 0)) if d_l else None,
                             width=width, rows=rows, stride_dest=sd, stride_src=ss, stride_alpha_a=saa,
                             stride_alpha_b=sab, weight=weight, softness=soft, step=step, op=op,
                             scalar_only=scalar_only, batch=1)
    ctx.call("mltcuda_composite", desc)
    return ctx.get(d_dest, dest.shape), (ctx.get(d_aa, aa.shape) if aa is not None else None)


@pytest.mark.parametrize("seed", range(40))
def test_composite_rows_random(ctx, seed):
    """Random sub-rectangles, offsets (odd and even), alpha/luma presence, weights, operators."""
    rng = np.random.default_rng(seed)
    wd, hd = int(rng.integers(8, 200)), int(rng.integers(2, 20))
    ws, hs = int(rng.integers(8, 200)), int(rng.integers(2, 20))
    width = int(rng.integers(1, min(wd, ws) + 1))
    rows = int(rng.integers(1, min(hd, hs) + 1))
    x, y = int(rng.integers(0, wd - width + 1)), int(rng.integers(0, hd - rows + 1))
    xs, ys = int(rng.integers(0, ws - width + 1)), int(rng.integers(0, hs - rows + 1))
    dest, src = rnd(rng, wd * hd * 2), rnd(rng, ws * hs * 2)
    ab = rnd(rng, ws * hs) if rng.random() < 0.5 else None
    aa = rnd(rng, wd * hd) if rng.random() < 0.5 else None
    luma = rnd(rng, ws * hs, np.uint16) if rng.random() < 0.3 else None
    weight = int(rng.choice([65536, 65535, 65280, 32768, 0, int(rng.integers(0, 65537))]))
    soft, step = int(rng.integers(0, 65536)), int(rng.integers(0, 131072))
    op = int(rng.choice([0, 0, 0, 1, 2, 3]))
    scalar_only = int(rng.random() < 0.15)
    off = dict(dest=2 * (x + y * wd), src=2 * (xs + ys * ws), alpha_b=xs + ys * ws, alpha_a=x + y * wd, luma=xs + ys * ws)
    args = (off, width, rows, wd * 2, ws * 2, wd, ws, weight, soft, step, op, scalar_only)
    exp, exp_a = oracle.composite_rows(dest, src, ab, aa, luma, *args)
    got, got_a = run_composite_rows(ctx, dest, src, ab, aa, luma, *args)
    assert np.array_equal(got, exp)
    if aa is not None:
        assert np.array_equal(got_a, exp_a)


@pytest.mark.parametrize("cond", range(8))
def test_composite_sse2_cases_full_rows(ctx, cond):
    """The 8 cases of composite_line_yuv_sse2_simple on 1920-wide rows + a width%8 tail variant."""
    rng = np.random.default_rng(100 + cond)
    for w in (1920, 1923):
        h = 4
        dest, src = rnd(rng, w * h * 2), rnd(rng, w * h * 2)
        aa = rnd(rng, w * h) if cond & 4 else None
        ab = rnd(rng, w * h) if cond & 2 else None
        weight = 32768 if cond & 1 else 65536
        off = dict(dest=0, src=0)
        args = (off, w, h, w * 2, w * 2, w, w, weight, 0, 0)
        exp, exp_a = oracle.composite_rows(dest, src, ab, aa, None, *args)
        got, got_a = run_composite_rows(ctx, dest, src, ab, aa, None, *args)
        assert np.array_equal(got, exp)
        if aa is not None:
            assert np.array_equal(got_a, exp_a)


@pytest.mark.parametrize("seed", range(24))
def test_composite_setup_geometry(ctx, seed):
    """mltcuda_composite_setup (host clipping) + kernel vs the oracle's composite_yuv, incl. fields,
    negative positions, crops and odd x (chroma re-alignment)."""
    rng = np.random.default_rng(1000 + seed)
    wd, hd = int(rng.choice([64, 96, 130])), int(rng.integers(8, 40))
    ws, hs = int(rng.integers(8, 120)) & ~1, int(rng.integers(4, 40))
    # two rows of padding: the reference's field stepping / chroma re-alignment may touch up to one
    # row (dest) or one pixel (src) past the logical image (harmless in its power-of-two pool)
    dest, src = rnd(rng, wd * (hd + 2) * 2), rnd(rng, ws * (hs + 2) * 2)
    ab = rnd(rng, ws * (hs + 2)) if rng.random() < 0.5 else None
    aa = rnd(rng, wd * (hd + 2)) if rng.random() < 0.5 else None
    luma = rnd(rng, ws * (hs + 2), np.uint16) if rng.random() < 0.3 else None
    nw, nh = wd, hd
    item = [float(rng.integers(-20, wd)), float(rng.integers(-10, hd)), float(ws), float(hs),
            float(rng.choice([100.0, 50.0, 70.0, 33.3]))]
    x_src = int(rng.choice([0, 0, -int(rng.integers(0, ws // 2))]))
    y_src = int(rng.choice([0, 0, -int(rng.integers(0, hs // 2))]))
    gi = [nw, nh, ws, hs, x_src, y_src]
    field = int(rng.choice([-1, -1, 0, 1]))
    softness = float(rng.choice([0.0, 0.2]))
    exp, exp_a, rows = oracle.composite_yuv(dest, wd, hd, src, ws, hs, ab, aa, item, gi, field, luma, softness)
    d_dest, d_src = ctx.put(dest), ctx.put(src)
    d_ab = ctx.put(ab) if ab is not None else None
    d_aa = ctx.put(aa) if aa is not None else None
    d_l = ctx.put(luma) if luma is not None else None
    g = abi.CompositeGeometry(x=item[0], y=item[1], w=item[2], h=item[3], o=item[4], nw=nw, nh=nh, sw=ws, sh=hs,
                              x_src=x_src, y_src=y_src)
    desc = abi.CompositeDesc(batch=1)
    skip = ctx.lib.mltcuda_composite_setup(C.byref(desc), d_dest.ptr, wd, hd, d_src.ptr, ws, hs,
                                           d_ab.ptr if d_ab else None, d_aa.ptr if d_aa else None, C.byref(g),
                                           field, d_l.ptr if d_l else None, softness)
    if not skip:
        ctx.call("mltcuda_composite", desc)
    assert (rows == 0) == bool(skip) or rows == desc.rows
    assert np.array_equal(ctx.get(d_dest, dest.shape), exp)
    if aa is not None:
        assert np.array_equal(ctx.get(d_aa, aa.shape), exp_a)


# ------------------------------------------------------------------------------------------------
# K3 luma / dissolve

def make_map(rng, lw, lh):
    ramp = (np.arange(lw * lh, dtype=np.uint32) * 65535 // max(lw * lh - 1, 1)).astype(np.uint16)
    return np.where(rng.random(lw * lh) < 0.3, rnd(rng, lw * lh, np.uint16), ramp).astype(np.uint16)


@pytest.mark.parametrize("w,h,lw,lh", [(64, 8, 64, 8), (1920, 6, 720, 576), (100, 7, 33, 21), (70, 9, 720, 576)])
@pytest.mark.parametrize("field_order", [-1, 0, 1])
@pytest.mark.parametrize("mode", ["opaque", "translucent_over", "translucent_add", "translucent_invert", "auto"])
def test_luma_yuv422(ctx, w, h, lw, lh, field_order, mode):
    rng = np.random.default_rng(w + h * 3 + lw + (field_order + 2) * 17)
    dest, src, lmap = rnd(rng, w * h * 2), rnd(rng, w * h * 2), make_map(rng, lw, lh)
    alpha = mode != "opaque"
    ad = rnd(rng, w * h) if alpha else None
    as_ = rnd(rng, w * h) if alpha else None
    if mode == "auto":
        ad[:] = 0xff
        as_[:] = 0xff
        as_[rng.integers(0, w * h)] = 0x10 if rng.random() < 0.5 else 0xff
    trans = {"opaque": 0, "auto": -1}.get(mode, 1)
    invert, over = int(mode == "translucent_invert"), int(mode != "translucent_add")
    pos, delta, soft = float(rng.uniform(0, 1)), 0.004, float(rng.choice([0.0, 0.2, 0.7]))
    # the map row index uses luma_width*field_count: keep it inside the map
    if field_order >= 0:
        lmap = np.concatenate([lmap, lmap, lmap[:2 * lw]])
    exp, exp_ad, exp_as, _ = oracle.luma_yuv422(dest, src, ad, as_, lmap, lw, lh, w, h, w, h, w, h, pos, delta, soft,
                                                field_order, invert, over, trans)
    d_dest, d_src, d_map = ctx.put(dest), ctx.put(src), ctx.put(lmap)
    d_ad = ctx.put(ad) if alpha else None
    d_as = ctx.put(as_) if alpha else None
    desc = abi.LumaDesc(format=YUV422, dest=d_dest.ptr, src=d_src.ptr, alpha_dest=d_ad.ptr if d_ad else None,
                        alpha_src=d_as.ptr if d_as else None, map=d_map.ptr, luma_width=lw, luma_height=lh, width=w,
                        height=h, width_dest=w, height_dest=h, width_src=w, height_src=h, pos=pos, frame_delta=delta,
                        softness=soft, field_order=field_order, invert=invert, alpha_over=over, translucent=trans,
                        batch=1)
    ctx.call("mltcuda_luma", desc)
    got = ctx.get(d_dest, dest.shape)
    tol = 0 if mode == "opaque" else 1
    assert np.abs(got.astype(int) - exp.astype(int)).max() <= tol
    if alpha:
        assert np.abs(ctx.get(d_ad, ad.shape).astype(int) - exp_ad.astype(int)).max() <= tol
        assert np.abs(ctx.get(d_as, as_.shape).astype(int) - exp_as.astype(int)).max() <= tol


@pytest.mark.parametrize("bits", [8, 16])
@pytest.mark.parametrize("w,h", [(64, 8), (1920, 4), (37, 5)])
@pytest.mark.parametrize("mode", ["opaque", "translucent_over", "translucent_add", "auto"])
def test_luma_rgba(ctx, bits, w, h, mode):
    rng = np.random.default_rng(bits + w + h)
    dt = np.uint8 if bits == 8 else np.uint16
    full = 0xff if bits == 8 else 0xffff
    dest, src = rnd(rng, w * h * 4, dt), rnd(rng, w * h * 4, dt)
    lw, lh = 50, 30
    lmap = make_map(rng, lw, lh)
    if mode in ("opaque", "auto"):
        dest[3::4] = full
        src[3::4] = full
    if mode == "auto" and rng.random() < 0.7:
        src[3 + 4 * int(rng.integers(0, w * h))] = 3
    trans = {"opaque": 0, "auto": -1}.get(mode, 1)
    over = int(mode != "translucent_add")
    pos, soft = float(rng.uniform(0, 1)), 0.3
    exp, _ = oracle.luma_rgba(bits, dest, src, lmap, lw, lh, w, h, pos, 0.004, soft, -1, 0, over, trans)
    d_dest, d_src, d_map = ctx.put(dest), ctx.put(src), ctx.put(lmap)
    desc = abi.LumaDesc(format=RGBA if bits == 8 else RGBA64, dest=d_dest.ptr, src=d_src.ptr, map=d_map.ptr,
                        luma_width=lw, luma_height=lh, width=w, height=h, width_dest=w, height_dest=h, width_src=w,
                        height_src=h, pos=pos, frame_delta=0.004, softness=soft, field_order=-1, invert=0,
                        alpha_over=over, translucent=trans, batch=1)
    ctx.call("mltcuda_luma", desc)
    got = ctx.get(d_dest, dest.shape, dt)
    tol = 0 if mode == "opaque" else 1
    assert np.abs(got.astype(int) - exp.astype(int)).max() <= tol


@pytest.mark.parametrize("w,h", [(64, 8), (1923, 3)])
@pytest.mark.parametrize("alpha", ["none", "opaque", "translucent"])
@pytest.mark.parametrize("alpha_over", [0, 1])
def test_dissolve_yuv422(ctx, w, h, alpha, alpha_over):
    """dissolve_yuv422: integer composite_line_yuv path unless translucent && alpha_over."""
    rng = np.random.default_rng(w + h)
    dest, src = rnd(rng, w * h * 2), rnd(rng, w * h * 2)
    aa = ab = None
    if alpha != "none":
        aa, ab = np.full(w * h, 255, np.uint8), np.full(w * h, 255, np.uint8)
        if alpha == "translucent":
            aa, ab = rnd(rng, w * h), rnd(rng, w * h)
    weight = 0.37
    if alpha == "translucent" and alpha_over:
        exp, exp_a = oracle.dissolve_yuv422_float(dest, src, aa, ab, w, h, weight)
        tol = 1
    else:
        off = dict(dest=0, src=0)
        exp, exp_a = oracle.composite_rows(dest, src, ab, aa, None, off, w, h, w * 2, w * 2, w, w,
                                           int(np.float32(weight) * np.float32(65536)), 0, 0)
        tol = 0
    d_dest, d_src = ctx.put(dest), ctx.put(src)
    d_aa = ctx.put(aa) if aa is not None else None
    d_ab = ctx.put(ab) if ab is not None else None
    desc = abi.LumaDesc(format=YUV422, dest=d_dest.ptr, src=d_src.ptr, alpha_dest=d_aa.ptr if d_aa else None,
                        alpha_src=d_ab.ptr if d_ab else None, map=None, width=w, height=h, width_dest=w, height_dest=h,
                        width_src=w, height_src=h, pos=weight, field_order=-1, alpha_over=alpha_over, translucent=-1,
                        batch=1)
    ctx.call("mltcuda_luma", desc)
    assert np.abs(ctx.get(d_dest, dest.shape).astype(int) - exp.astype(int)).max() <= tol
    if aa is not None:
        assert np.abs(ctx.get(d_aa, aa.shape).astype(int) - exp_a.astype(int)).max() <= tol


@pytest.mark.parametrize("bits", [8, 16])
def test_dissolve_rgba(ctx, bits):
    rng = np.random.default_rng(bits)
    w, h = 101, 7
    dt = np.uint8 if bits == 8 else np.uint16
    dest, src = rnd(rng, w * h * 4, dt), rnd(rng, w * h * 4, dt)
    exp = oracle.dissolve_rgba(bits, dest, src, w, h, 0.61)
    d_dest, d_src = ctx.put(dest), ctx.put(src)
    desc = abi.LumaDesc(format=RGBA if bits == 8 else RGBA64, dest=d_dest.ptr, src=d_src.ptr, map=None, width=w,
                        height=h, width_dest=w, height_dest=h, width_src=w, height_src=h, pos=0.61, field_order=-1,
                        alpha_over=1, translucent=1, batch=1)
    ctx.call("mltcuda_luma", desc)
    assert np.abs(ctx.get(d_dest, dest.shape, dt).astype(int) - exp.astype(int)).max() <= 1


def test_scale_luma(ctx):
    rng = np.random.default_rng(3)
    src = rnd(rng, 720 * 576, np.uint16)
    for dw, dh, inv in [(1920, 1080, 0), (960, 540, 0xffff), (33, 17, 0)]:
        exp = oracle.scale_luma(src, 720, 576, dw, dh, inv)
        d_src, d_dst = ctx.put(src), ctx.put(np.zeros(dw * dh, np.uint16))
        abi.check(ctx.lib.mltcuda_scale_luma(ctx.h, d_dst.ptr, dw, dh, d_src.ptr, 720, 576, inv))
        assert np.array_equal(ctx.get(d_dst, exp.shape, np.uint16), exp)


# ------------------------------------------------------------------------------------------------
# K4 affine

def affine_case(rng, aw, ah, bw, bh, angle, scale, interp, mix=1.0, b_alpha=0):
    c, s = np.cos(np.radians(angle)), np.sin(np.radians(angle))
    m = np.array([[c * scale, -s * scale, 0.0], [s * scale, c * scale, 0.0], [0, 0, 1.0]])
    minima, xmax, ymax = 0.0, bw - 1.0, bh - 1.0
    if interp == 0:
        minima, xmax, ymax = -0.5, xmax + 0.49, ymax + 0.49
    elif interp == 2:
        minima = -1.0
    return dict(m=m, lower_x=-float(int(aw * 0.45)), lower_y=-float(ah // 2), dz=1.0, mix=mix,
                x_offset=bw / 2.0, y_offset=bh / 2.0, minima=minima, xmax=xmax, ymax=ymax, b_alpha=b_alpha,
                interp=interp)


def run_affine(ctx, bits, a, b, aw, ah, bw, bh, k):
    d_a, d_b = ctx.put(a), ctx.put(b)
    desc = abi.AffineDesc(format=RGBA if bits == 8 else RGBA64, a_image=d_a.ptr, b_image=d_b.ptr, a_width=aw,
                          a_height=ah, b_width=bw, b_height=bh, lower_x=k["lower_x"], lower_y=k["lower_y"],
                          dz=k["dz"], mix=k["mix"], x_offset=k["x_offset"], y_offset=k["y_offset"],
                          minima=k["minima"], xmax=k["xmax"], ymax=k["ymax"], b_alpha=k["b_alpha"],
                          interp=k["interp"], batch=1)
    for i in range(3):
        for j in range(3):
            desc.matrix[i][j] = float(k["m"][i][j])
    ctx.call("mltcuda_affine", desc)
    return ctx.get(d_a, a.shape, a.dtype)


@pytest.mark.parametrize("bits", [8, 16])
@pytest.mark.parametrize("interp", [0, 1, 2])
@pytest.mark.parametrize("angle,scale", [(0.0, 1.0), (17.0, 1.3), (90.0, 0.5), (-33.0, 2.2)])
@pytest.mark.parametrize("mix,b_alpha", [(1.0, 0), (0.6, 0), (1.0, 1)])
def test_affine(ctx, bits, interp, angle, scale, mix, b_alpha):
    rng = np.random.default_rng(bits * 7 + interp * 31 + int(angle))
    aw, ah, bw, bh = 160, 90, 64, 48
    dt = np.uint8 if bits == 8 else np.uint16
    a, b = rnd(rng, aw * ah * 4, dt), rnd(rng, bw * bh * 4, dt)
    b[3::4][rng.random(bw * bh) < 0.1] = 0  # zero source alpha: 0/0 -> NaN -> stored as 0 (App. A5)
    a[3::4][rng.random(aw * ah) < 0.1] = 0
    k = affine_case(rng, aw, ah, bw, bh, angle, scale, interp, mix, b_alpha)
    exp, covered = oracle.affine(bits, a, b, aw, ah, bw, bh, k["m"], k["lower_x"], k["lower_y"], k["dz"], k["mix"],
                                 k["x_offset"], k["y_offset"], k["minima"], k["xmax"], k["ymax"], b_alpha, interp)
    assert covered > 0
    got = run_affine(ctx, bits, a, b, aw, ah, bw, bh, k)
    diff = np.abs(got.astype(int) - exp.astype(int))
    # The oracle runs in the operation order of the reference's Release build (pinned bit-exact
    # against the compiled samplers in test_oracle_pinning.py) and the kernel follows the same
    # order, so the result is bit-exact, not merely within the +-1 LSB the spec allows.  (The oracle
    # steps x/y with repeated ++ in double, the kernel adds the index once: identical for these
    # half-integer origins.)
    assert diff.max() == 0, f"{(diff > 0).sum()} bytes differ, max {diff.max()} at {np.flatnonzero(diff > 0)[:5]}"


# ------------------------------------------------------------------------------------------------
# producers / normalisers

@pytest.mark.parametrize("w,h,pos", [(1920, 1080, 0), (1920, 1080, 249), (1536, 864, 10), (64, 3, 5), (37, 3, 100000)])
def test_noise(ctx, w, h, pos):
    exp = oracle.noise(w, h, pos)
    d = ctx.put(np.zeros(w * h * 2, np.uint8))
    abi.check(ctx.lib.mltcuda_noise(ctx.h, d.ptr, w, h, pos, 1, 0))
    assert np.array_equal(ctx.get(d, exp.shape), exp)


def test_noise_batch_positions(ctx):
    w, h, n = 128, 16, 4
    d = ctx.put(np.zeros(n * w * h * 2, np.uint8))
    abi.check(ctx.lib.mltcuda_noise(ctx.h, d.ptr, w, h, 7, n, w * h * 2))
    got = ctx.get(d, (n, w * h * 2))
    for i in range(n):
        assert np.array_equal(got[i], oracle.noise(w, h, 7 + i))


@pytest.mark.parametrize("fmt", [YUV422, RGBA, RGBA64])
def test_colour(ctx, fmt):
    w, h = 65, 9
    exp, exp_a = oracle.colour(fmt, w, h, 0x30, 0x60, 0xc0, 0xa0)
    d, da = ctx.put(np.zeros(exp.size, np.uint8)), ctx.put(np.zeros(w * h, np.uint8))
    abi.check(ctx.lib.mltcuda_colour(ctx.h, d.ptr, da.ptr, fmt, w, h, 0x30, 0x60, 0xc0, 0xa0))
    assert np.array_equal(ctx.get(d, exp.shape), exp)
    if fmt == YUV422:
        assert np.array_equal(ctx.get(da, exp_a.shape), exp_a)


@pytest.mark.parametrize("iw,ih,ow,oh", [(3840, 216, 1536, 86), (1920, 108, 960, 54), (100, 60, 160, 90), (64, 32, 30, 17)])
def test_rescale(ctx, iw, ih, ow, oh):
    rng = np.random.default_rng(iw + ow)
    src = rnd(rng, iw * ih * 2)
    init = rnd(rng, ow * oh * 2)
    exp = oracle.rescale_yuv422(src, iw, ih, ow, oh, dst_init=init)
    d_src, d_dst = ctx.put(src), ctx.put(init)
    abi.check(ctx.lib.mltcuda_rescale_yuv422(ctx.h, d_dst.ptr, ow, oh, d_src.ptr, iw, ih, 1, 0, 0))
    assert np.array_equal(ctx.get(d_dst, exp.shape), exp)


@pytest.mark.parametrize("w,h,alpha", [(1920, 1080, False), (1536, 864, True), (642, 37, True), (2, 3, False), (10, 5, True)])
def test_mirror_flip(ctx, w, h, alpha):
    """core mirror:flip (filter_mirror.c:247-294): rows reversed pair-wise with the lumas of a pair swapped."""
    rng = np.random.default_rng(w * 31 + h)
    src = rng.integers(0, 256, w * h * 2).astype(np.uint8)
    al = rng.integers(0, 256, w * h).astype(np.uint8) if alpha else None
    exp = np.ascontiguousarray(src.reshape(h, w // 2, 4)[:, ::-1, :][:, :, [2, 1, 0, 3]]).reshape(-1)
    d_src, d_dst = ctx.put(src), ctx.alloc(w * h * 2)
    d_al = ctx.put(al) if alpha else None
    d_alo = ctx.alloc(w * h) if alpha else None
    abi.check(ctx.lib.mltcuda_mirror_flip(ctx.h, d_dst.ptr, d_alo.ptr if alpha else None, d_src.ptr, d_al.ptr if alpha else None, w, h))
    assert np.array_equal(ctx.get(d_dst, (w * h * 2,)), exp)
    if alpha:
        assert np.array_equal(ctx.get(d_alo, (w * h,)), np.ascontiguousarray(al.reshape(h, w)[:, ::-1]).reshape(-1))
    assert ctx.lib.mltcuda_mirror_flip(ctx.h, d_dst.ptr, None, d_src.ptr, None, 7, 4) == abi.E_UNSUPPORTED


def _ref_resize_pad(src, iw, ih, ow, oh, bpp, fill_px, yuv):
    """resize_image (filter_resize.c:65-137): fill, then the input rows copied at the centred offset."""
    out = np.tile(np.array(fill_px, np.uint8), ow * oh).reshape(oh, ow * bpp)
    off_x = (ow - iw) // 2 * bpp
    if yuv:
        off_x -= off_x % 4
    off_y = (oh - ih) // 2
    out[off_y:off_y + ih, off_x:off_x + iw * bpp] = src.reshape(ih, iw * bpp)
    return out.reshape(-1)


@pytest.mark.parametrize("fmt,bpp,iw,ih,ow,oh,alpha", [(3, 2, 1440, 1080, 1920, 1080, 0), (3, 2, 630, 355, 632, 355, 7),
                                                      (2, 4, 100, 50, 128, 72, 200), (10, 8, 64, 40, 96, 54, 255),
                                                      (1, 3, 33, 20, 40, 30, 0), (0, 1, 101, 37, 128, 41, 99)])
def test_resize_pad(ctx, fmt, bpp, iw, ih, ow, oh, alpha):
    rng = np.random.default_rng(iw * 7 + ow)
    src = rng.integers(0, 256, iw * ih * bpp).astype(np.uint8)
    fill = {3: [16, 128], 2: [0, 0, 0, alpha], 10: [0, 0, 0, 0, 0, 0, 0, alpha], 1: [0, 0, 0], 0: [alpha]}[fmt]
    if fmt == 0:  # resize_alpha (:30-63): offset_x rounded down to even
        exp = np.full((oh, ow), alpha, np.uint8)
        ox, oy = (ow - iw) // 2, (oh - ih) // 2
        ox -= ox % 2
        exp[oy:oy + ih, ox:ox + iw] = src.reshape(ih, iw)
        exp = exp.reshape(-1)
    else:
        exp = _ref_resize_pad(src, iw, ih, ow, oh, bpp, fill, fmt == 3)
    d_src, d_dst = ctx.put(src), ctx.alloc(ow * oh * bpp)
    abi.check(ctx.lib.mltcuda_resize_pad(ctx.h, d_dst.ptr, ow, oh, d_src.ptr, iw, ih, fmt, alpha))
    assert np.array_equal(ctx.get(d_dst, (ow * oh * bpp,)), exp)


@pytest.mark.parametrize("iw,ih,ow,oh", [(1920, 1080, 576, 324), (720, 576, 1920, 1080), (101, 57, 64, 33)])
def test_rescale_alpha(ctx, iw, ih, ow, oh):
    """scale_alpha (filter_rescale.c:122-147): nearest neighbour in 16.16 fixed point, sampled at pixel centres."""
    rng = np.random.default_rng(iw + oh)
    src = rng.integers(0, 256, iw * ih).astype(np.uint8)
    ox, oy = (iw << 16) // ow, (ih << 16) // oh
    ys = ((oy >> 1) + np.arange(oh) * oy) >> 16
    xs = ((ox >> 1) + np.arange(ow) * ox) >> 16
    exp = src.reshape(ih, iw)[ys][:, xs].reshape(-1)
    d_src, d_dst = ctx.put(src), ctx.alloc(ow * oh)
    abi.check(ctx.lib.mltcuda_rescale_alpha(ctx.h, d_dst.ptr, ow, oh, d_src.ptr, iw, ih))
    assert np.array_equal(ctx.get(d_dst, (ow * oh,)), exp)
