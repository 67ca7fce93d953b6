"""ctypes access to the UNMODIFIED reference compiled into oracle/_ref (test infrastructure).

Only exported hot-path symbols are called: composite_line_yuv (libmltcore.so) and the
interp*_b32/_b64 samplers (libmltplus.so).  `available()` is False when oracle/_ref was not built.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
_libs = None
vp = C.c_void_p


def available():
    return os.path.exists(os.path.join(REF, "lib", "mlt", "libmltcore.so"))


def libs():
    global _libs
    if _libs is None:
        mlt = C.CDLL(os.path.join(REF, "lib", "libmlt-7.so"), mode=C.RTLD_GLOBAL)
        core = C.CDLL(os.path.join(REF, "lib", "mlt", "libmltcore.so"))
        plus = C.CDLL(os.path.join(REF, "lib", "mlt", "libmltplus.so"))
        _libs = (mlt, core, plus)
    return _libs


def env():
    e = dict(os.environ)
    e["MLT_REPOSITORY"] = os.path.join(REF, "lib", "mlt")
    e["MLT_DATA"] = os.path.join(REF, "share", "mlt")
    e["MLT_PROFILES_PATH"] = os.path.join(REF, "share", "mlt", "profiles")
    e["LD_LIBRARY_PATH"] = os.path.join(REF, "lib") + ":" + e.get("LD_LIBRARY_PATH", "")
    return e


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


def composite_line_yuv(dest, src, width, alpha_b, alpha_a, weight, luma, soft, step):
    d = dest.copy()
    aa = None if alpha_a is None else alpha_a.copy()
    libs()[1].composite_line_yuv(_p(d), _p(src), C.c_int(width), _p(alpha_b), _p(aa), C.c_int(weight), _p(luma),
                                 C.c_int(soft), C.c_uint32(step))
    return d, aa


def interp(bits, kind, s, w, h, x, y, o, d, is_atop):
    name = "interp%s_b%d" % (["NN", "BL", "BC"][kind], 32 if bits == 8 else 64)
    dd = d.copy()
    getattr(libs()[2], name)(_p(s), C.c_int(w), C.c_int(h), C.c_float(x), C.c_float(y), C.c_float(o), _p(dd),
                             C.c_int(is_atop))
    return dd
