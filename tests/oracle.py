"""ctypes binding of oracle/liboracle.so (the plain-C CPU restatement) -- the parity checker.

Test infrastructure only.  numpy in, numpy out; inputs that the reference modifies in place are
copied first so callers keep their originals.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")

_lib = None
vp = C.c_void_p


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "mlt_oracle.c")
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(LIB)
        _lib.orc_affine.restype = C.c_long
        # oracle of record = the Release x86-64 build: float paths in the operation order
        # that build executes; test_oracle_pinning.py checks both orders against the compiled .so
        _lib.orc_set_compiled_order(C.c_int(1))
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


def composite_line(dest, src, width, alpha_b=None, alpha_a=None, weight=65536, luma=None, soft=0, step=0, op=0,
                   scalar_only=0):
    d = dest.copy()
    aa = None if alpha_a is None else alpha_a.copy()
    lib().orc_composite_line_yuv(_p(d), _p(src), C.c_int(width), _p(alpha_b), _p(aa), C.c_int(weight), _p(luma),
                                 C.c_int(soft), C.c_uint32(step), C.c_int(op), C.c_int(scalar_only))
    return d, aa


def composite_rows(dest, src, alpha_b, alpha_a, luma, off, width, rows, stride_dest, stride_src, stride_aa,
                   stride_ab, weight, soft, step, op=0, scalar_only=0):
    """off = dict(dest=, src=, alpha_b=, alpha_a=, luma=) element offsets (bytes for images)."""
    d = dest.copy()
    aa = None if alpha_a is None else alpha_a.copy()

    def at(a, o, itemsize=1):
        return None if a is None else C.c_void_p(a.ctypes.data + o * itemsize)

    lib().orc_composite_rows(at(d, off["dest"]), at(src, off["src"]), at(alpha_b, off.get("alpha_b", 0)),
                             at(aa, off.get("alpha_a", 0)), at(luma, off.get("luma", 0), 2), C.c_int(width),
                             C.c_int(rows), C.c_int(stride_dest), C.c_int(stride_src), C.c_int(stride_aa),
                             C.c_int(stride_ab), C.c_int(weight), C.c_int(soft), C.c_uint32(step), C.c_int(op),
                             C.c_int(scalar_only))
    return d, aa


def composite_yuv(dest, wd, hd, src, ws, hs, alpha_b, alpha_a, item, gi, field=-1, luma=None, softness=0.0, op=0,
                  scalar_only=0):
    d = dest.copy()
    aa = None if alpha_a is None else alpha_a.copy()
    it = (C.c_double * 5)(*item)
    g = (C.c_int * 6)(*gi)
    rows = lib().orc_composite_yuv(_p(d), wd, hd, _p(src), ws, hs, _p(alpha_b), _p(aa), it, g, C.c_int(field),
                                   _p(luma), C.c_double(softness), C.c_int(op), C.c_int(scalar_only))
    return d, aa, rows


BPP = {1: 3, 2: 4, 3: 2, 10: 8}


def image_bytes(fmt, w, h):
    if fmt == 4:
        return w * h + 2 * ((w // 2) * (h // 2))
    return w * h * BPP[fmt]


def convert(sf, df, w, h, src, src_alpha=None, want_alpha=False, dst_init=None):
    dst = np.zeros(image_bytes(df, w, h), np.uint8) if dst_init is None else dst_init.copy()
    da = np.zeros(w * h, np.uint8) if want_alpha else None
    rc = lib().orc_convert(sf, df, w, h, _p(src), _p(dst), _p(src_alpha), _p(da))
    return rc, dst, da


def luma_yuv422(dest, src, alpha_dest, alpha_src, lmap, lw, lh, width, height, wd, hd, ws, hs, pos, frame_delta,
                softness, field_order=-1, invert=0, alpha_over=0, translucent=-1):
    d = dest.copy()
    ad = None if alpha_dest is None else alpha_dest.copy()
    as_ = None if alpha_src is None else alpha_src.copy()
    flag = lib().orc_luma_yuv422(_p(d), _p(src), _p(ad), _p(as_), _p(lmap), lw, lh, width, height, wd, hd, ws, hs,
                                 C.c_float(pos), C.c_float(frame_delta), C.c_float(softness), C.c_int(field_order),
                                 invert, alpha_over, C.c_int(translucent))
    return d, ad, as_, flag


def luma_rgba(bits, dest, src, lmap, lw, lh, w, h, pos, frame_delta, softness, field_order=-1, invert=0,
              alpha_over=0, translucent=-1):
    d = dest.copy()
    flag = lib().orc_luma_rgba(bits, _p(d), _p(src), _p(lmap), lw, lh, w, h, C.c_float(pos), C.c_float(frame_delta),
                               C.c_float(softness), C.c_int(field_order), invert, alpha_over, C.c_int(translucent))
    return d, flag


def dissolve_yuv422_float(dest, src, alpha_a, alpha_b, w, h, weight):
    d = dest.copy()
    aa = None if alpha_a is None else alpha_a.copy()
    lib().orc_dissolve_yuv422_float(_p(d), _p(src), _p(aa), _p(alpha_b), w, h, C.c_float(weight))
    return d, aa


def dissolve_rgba(bits, dest, src, w, h, weight):
    d = dest.copy()
    lib().orc_dissolve_rgba(bits, _p(d), _p(src), w, h, C.c_float(weight))
    return d


def set_compiled_order(on):
    """1 = operation order of the reference's -ffast-math Release build (see mlt_oracle.c)."""
    lib().orc_set_compiled_order(C.c_int(1 if on else 0))


def interp_px(bits, kind, s, w, h, x, y, o, d, is_atop):
    dd = d.copy()
    lib().orc_interp_px(bits, kind, _p(s), w, h, C.c_float(x), C.c_float(y), C.c_float(o), _p(dd), is_atop)
    return dd


def affine(bits, a_image, b_image, aw, ah, bw, bh, m, lower_x, lower_y, dz, mix, x_offset, y_offset, minima, xmax,
           ymax, b_alpha, interp):
    a = a_image.copy()
    mm = (C.c_double * 9)(*[float(v) for v in np.asarray(m).reshape(9)])
    cd = C.c_double
    covered = lib().orc_affine(bits, _p(a), _p(b_image), aw, ah, bw, bh, mm, cd(lower_x), cd(lower_y), cd(dz), cd(mix),
                               cd(x_offset), cd(y_offset), cd(minima), cd(xmax), cd(ymax), b_alpha, interp)
    return a, covered


def noise(w, h, position):
    out = np.zeros(w * h * 2, np.uint8)
    lib().orc_noise(_p(out), w, h, position)
    return out


def rescale_yuv422(src, iw, ih, ow, oh, dst_init=None):
    out = np.zeros(ow * oh * 2, np.uint8) if dst_init is None else dst_init.copy()
    lib().orc_rescale_yuv422(_p(out), ow, oh, _p(src), iw, ih)
    return out


def scale_luma(src, sw, sh, dw, dh, invert_xor=0):
    out = np.zeros(dw * dh, np.uint16)
    lib().orc_scale_luma(_p(out), dw, dh, _p(src), sw, sh, invert_xor)
    return out


def colour(fmt, w, h, r, g, b, a):
    out = np.zeros(image_bytes(fmt, w, h), np.uint8)
    al = np.zeros(w * h, np.uint8)
    lib().orc_colour(_p(out), _p(al), fmt, w, h, C.c_uint8(r), C.c_uint8(g), C.c_uint8(b), C.c_uint8(a))
    return out, al
