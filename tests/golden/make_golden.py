#!/usr/bin/env python
"""Generates tests/golden/*.npz / *.json from the UNMODIFIED reference compiled into oracle/_ref
(oracle/Makefile.ref).  Run in a container that has /root/reference:  python tests/golden/make_golden.py
The fixtures are what the parity tests fall back on when oracle/_ref is absent; they are small
(seeded inputs are regenerated, only the reference's outputs are stored)."""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import mltenv  # noqa: E402
import refmlt  # noqa: E402

SEED = 20261104


def line_cases():
    """(name, width, weight, has_alpha_b, has_alpha_a, has_luma, soft, step)"""
    out = []
    for w in (1, 7, 8, 9, 16, 33, 255, 300):
        for weight in (65536, 32768, 45875, 0, 255):
            for ab in (0, 1):
                for aa in (0, 1):
                    out.append((f"w{w}_m{weight}_b{ab}_a{aa}", w, weight, ab, aa, 0, 0, 0))
        out.append((f"w{w}_luma", w, 30000, 1, 1, 1, 13107, 40000))
        out.append((f"w{w}_luma_noalpha", w, 30000, 0, 0, 1, 6553, 20000))
    return out


def line_inputs(rng, w, ab, aa, lum):
    dest = rng.integers(0, 256, 2 * w).astype(np.uint8)
    src = rng.integers(0, 256, 2 * w).astype(np.uint8)
    alpha_b = rng.integers(0, 256, w).astype(np.uint8) if ab else None
    alpha_a = rng.integers(0, 256, w).astype(np.uint8) if aa else None
    luma = rng.integers(0, 65536, w).astype(np.uint16) if lum else None
    return dest, src, alpha_b, alpha_a, luma


def interp_inputs(rng, bits):
    w = h = 9
    dt, hi = (np.uint8, 256) if bits == 8 else (np.uint16, 65536)
    s = rng.integers(0, hi, w * h * 4).astype(dt)
    d = rng.integers(0, hi, 4).astype(dt)
    x, y = float(rng.uniform(0.01, w - 1.0)), float(rng.uniform(0.01, h - 1.0))
    o = float(rng.choice([1.0, 0.8, 0.5, 0.33]))
    return s, d, w, h, x, y, o, int(rng.integers(0, 2))


FRAMES = {
    "cfg1a_blit": ("yuv422", 10, ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition", "composite", "a_track=0", "b_track=1"]),
    "cfg1b_blend": ("yuv422", 10, ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition", "composite", "a_track=0", "b_track=1", "geometry=0/0:100%x100%:50%"]),
    "cfg1c_subrect": ("yuv422", 10, ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition", "composite", "a_track=0", "b_track=1", "geometry=25%/25%:50%x50%:50%"]),
    "cfg1d_alpha_b": ("yuv422", 10, ["noise:", "out=249", "-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=1", "geometry=10%/10%:50%x50%:70%"]),
    "cfg2_luma_wipe": ("yuv422", 100, ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition", "luma:" + mltenv.LUMA_PGM, "softness=0.2", "in=0", "out=249", "a_track=0", "b_track=1"]),
    "cfg2_dissolve": ("yuv422", 100, ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition", "luma", "in=0", "out=249", "a_track=0", "b_track=1"]),
    "cfg3_affine_rgba": ("rgba", 100, ["noise:", "out=249", "-track", "noise:", "out=249", "meta.media.width=1920", "meta.media.height=1080", "-attach", "mirror:flip", "-transition", "affine", "rect=0=10%/10%:50%x50%:100%; 249=30%/30%:70%x70%:60%", "rotate_x=1", "in=0", "out=249", "a_track=0", "b_track=1"]),
    "cfg3_affine_yuv422": ("yuv422", 100, ["noise:", "out=249", "-track", "noise:", "out=249", "meta.media.width=1920", "meta.media.height=1080", "-attach", "mirror:flip", "-transition", "affine", "rect=0=10%/10%:50%x50%:100%; 249=30%/30%:70%x70%:60%", "rotate_x=1", "in=0", "out=249", "a_track=0", "b_track=1"]),
    "cfg4_1080p_f0": ("yuv422", 0, None),
    "cfg4_1080p_f37": ("yuv422", 37, None),
    # the full BASELINE size: cfg4 on uhd_2160p_60 (4th tuple element = profile)
    "cfg4_2160p_f300": ("yuv422", 300, None, "uhd_2160p_60"),
    "cfg4_2160p_f599": ("yuv422", 599, None, "uhd_2160p_60"),
}
CFG4 = ["noise:", "out=599", "-track", "noise:", "out=599", "-attach", "mirror:flip", "-transition", "composite", "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-track", "colour:0x3060c0a0", "out=599", "-transition", "affine", "a_track=0", "b_track=2", "rect=0=20%/20%:50%x50%:90%; 599=30%/25%:60%x60%:70%", "rotate_x=0.5", "-track", "noise:", "out=599", "-transition", "composite", "a_track=0", "b_track=3", "geometry=55%/55%:40%x40%:100%"]


def main():
    assert refmlt.available() and mltenv.have_reference(), "build oracle/_ref first (make -C oracle)"
    rng = np.random.default_rng(SEED)
    lines = {}
    for name, w, weight, ab, aa, lum, soft, step in line_cases():
        dest, src, alpha_b, alpha_a, luma = line_inputs(rng, w, ab, aa, lum)
        d, a = refmlt.composite_line_yuv(dest, src, w, alpha_b, alpha_a, weight, luma, soft, step)
        lines[name + "/dest"] = d
        if a is not None:
            lines[name + "/alpha_a"] = a
    np.savez_compressed(os.path.join(HERE, "composite_line_yuv.npz"), **lines)

    rng = np.random.default_rng(SEED + 1)
    interp = {}
    for bits in (8, 16):
        for kind in (0, 1, 2):
            outs = []
            for _ in range(400):
                s, d, w, h, x, y, o, atop = interp_inputs(rng, bits)
                outs.append(refmlt.interp(bits, kind, s, w, h, x, y, o, d, atop))
            interp[f"b{bits}_k{kind}"] = np.stack(outs)
    np.savez_compressed(os.path.join(HERE, "interp.npz"), **interp)

    hashes = {}
    td = tempfile.mkdtemp()
    for name, spec in FRAMES.items():
        fmt, frame, args = spec[:3]
        profile = spec[3] if len(spec) > 3 else "atsc_1080p_25"
        args = args or CFG4
        out = subprocess.check_output([mltenv.DUMP, profile, fmt, os.path.join(td, "g"), str(frame), "--"] + args,
                                      env=mltenv.env(), stderr=subprocess.DEVNULL).decode().split()
        hashes[name] = {"profile": profile, "format": fmt, "frame": frame, "width": int(out[1]),
                        "height": int(out[2]), "image_fnv1a64": out[4], "alpha_fnv1a64": out[5],
                        "args": [a.replace(mltenv.LUMA_PGM, "@LUMA1_PGM@") for a in args]}
    json.dump({"seed": SEED, "reference": "MLT 7.41.0, oracle/Makefile.ref Release flags, x86-64", "frames": hashes},
              open(os.path.join(HERE, "frames.json"), "w"), indent=1)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
