"""Pins the oracle chain + the host-side parameter derivation (mlt_b200/workloads.py) against
frames rendered by the UNMODIFIED reference (oracle/_ref, built by oracle/Makefile.ref) through its
public API (oracle/mlt_dump.c).  CPU only.  Skipped when oracle/_ref is not built."""
import os
import subprocess

import numpy as np
import pytest

import oracle
import pipeline
import refmlt

pytestmark = pytest.mark.skipif(not refmlt.available(), reason="oracle/_ref not built")

DUMP = os.path.join(refmlt.REF, "bin", "mlt_dump")
LUMA_PGM = os.path.join(refmlt.REF, "share", "mlt", "demo", "luma1.pgm")


def dump(tmp_path, profile, fmt, frame, args):
    prefix = str(tmp_path / "f")
    out = subprocess.check_output([DUMP, profile, fmt, prefix, str(frame), "--"] + args, env=refmlt.env(),
                                  stderr=subprocess.DEVNULL).decode().split()
    img = np.fromfile(f"{prefix}.{frame}.img", np.uint8)
    ap = f"{prefix}.{frame}.alpha"
    alpha = np.fromfile(ap, np.uint8) if os.path.exists(ap) else None
    return img, alpha, out


CFG4_ARGS = ["noise:", "out=599", "-track", "noise:", "out=599", "-attach", "mirror:flip", "-transition", "composite",
             "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-track", "colour:0x3060c0a0", "out=599",
             "-transition", "affine", "a_track=0", "b_track=2",
             "rect=0=20%/20%:50%x50%:90%; 599=30%/25%:60%x60%:70%", "rotate_x=0.5", "-track", "noise:", "out=599",
             "-transition", "composite", "a_track=0", "b_track=3", "geometry=55%/55%:40%x40%:100%"]


@pytest.mark.parametrize("frame", [0, 37])
def test_cfg4_oracle_chain_matches_reference_1080p(tmp_path, frame):
    W, H = 1920, 1080
    img, alpha, _ = dump(tmp_path, "atsc_1080p_25", "yuv422", frame, CFG4_ARGS)
    res = pipeline.oracle_chain(pipeline.inputs(frame, W, H))
    diff = np.abs(img.astype(int) - res["out"].astype(int))
    # composites and conversions are integer-exact; only pixels the float affine touched may differ
    assert diff.max() <= 2, diff.max()
    assert (diff > 0).mean() < 0.01
    assert alpha is not None and np.abs(alpha.astype(int) - res["alpha"].astype(int)).max() <= 1


def test_cfg1_composite_blend_matches_reference(tmp_path):
    W, H = 1920, 1080
    img, _, out = dump(tmp_path, "atsc_1080p_25", "yuv422", 10,
                       ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition",
                        "composite", "a_track=0", "b_track=1", "geometry=0/0:100%x100%:50%"])
    assert out[4] == "13c489d896346a6e"
    a = oracle.noise(W, H, 10)
    b = pipeline.flip_yuv422(oracle.noise(W, H, 10), W, H)
    exp, _, _ = oracle.composite_yuv(a, W, H, b, W, H, None, None, [0, 0, W, H, 50.0], [W, H, W, H, 0, 0])
    assert np.array_equal(img, exp)


def test_cfg2_luma_wipe_matches_reference(tmp_path):
    W, H = 1920, 1080
    frame = 100
    img, _, out = dump(tmp_path, "atsc_1080p_25", "yuv422", frame,
                       ["noise:", "out=249", "-track", "noise:", "out=249", "-attach", "mirror:flip", "-transition",
                        f"luma:{LUMA_PGM}", "softness=0.2", "in=0", "out=249", "a_track=0", "b_track=1"])
    assert out[4] == "2a9ef4d0fdc282ef"
    a = oracle.noise(W, H, frame)
    b = pipeline.flip_yuv422(oracle.noise(W, H, frame), W, H)
    pgm = open(LUMA_PGM, "rb").read()
    assert pgm[:2] == b"P5"
    lmap = (np.frombuffer(pgm[-720 * 576:], np.uint8).astype(np.uint16) << 8)  # mlt_luma_map.c:401-406
    pos = np.float32(frame / 250.0)  # mlt_transition_get_progress: (position-in)/(out-in+1) (mlt_transition.c:251-275)
    delta = np.float32(1 / 250.0)
    exp, _, _, _ = oracle.luma_yuv422(a, b, None, None, lmap, 720, 576, W, H, W, H, W, H, float(pos), float(delta),
                                      0.2, -1, 0, 0, 0)
    assert np.array_equal(img, exp)
