"""The MLT module (libmltcuda.so) inside the reference's own melt / libmlt-7 (oracle/_ref).

CPU tier: the module loads through mlt_repository, registers the reference's service names plus the
cuda.* aliases, serves its metadata, and fails loudly without a GPU.
GPU tier: whole frames rendered through the public MLT API with the stock services (override off)
and with the CUDA services (override on) from the same melt command line must agree -- bit-exact
for composite / luma / imageconvert, +-1 LSB (rgba) for affine.
"""
import os
import subprocess

import numpy as np
import pytest

import mltenv

pytestmark = pytest.mark.skipif(not (mltenv.have_reference() and mltenv.have_module()),
                                reason="oracle/_ref or libmltcuda.so not built")

NOISE_B = ["-track", "noise:", "out=249", "-attach", "mirror:flip"]
A = ["noise:", "out=249"]


def melt_query(tmp_path, what):
    e = mltenv.env(str(tmp_path), cuda=True)
    return subprocess.run([mltenv.MELT, "-query", what], env=e, capture_output=True, text=True).stdout


def test_module_registers_reference_names_and_aliases(tmp_path):
    out = melt_query(tmp_path, "transitions")
    for n in ("composite", "luma", "affine", "cuda.composite", "cuda.luma", "cuda.affine"):
        assert f"- {n}\n" in out
    out = melt_query(tmp_path, "filters")
    assert "- imageconvert\n" in out and "- cuda.imageconvert\n" in out
    # the reference names now resolve to the CUDA metadata
    assert "(CUDA)" in melt_query(tmp_path, "transition=composite")
    assert "(CUDA)" in melt_query(tmp_path, "transition=cuda.affine")


def test_metadata_lists_every_parameter_of_the_reference_services(tmp_path):
    """melt -query of the CUDA services shows at least the parameters the stock services document (SURVEY 8(b):
    same property names), parsed from the stock YAML shipped with the reference build."""
    import yaml
    for kind, name, mod in (("transition", "composite", "core"), ("transition", "luma", "core"), ("transition", "affine", "plus")):
        ref = yaml.safe_load(open(os.path.join(mltenv.REF, "share", "mlt", mod, f"{kind}_{name}.yml")))
        ours = yaml.safe_load(melt_query(tmp_path, f"{kind}={name}").replace("...", ""))
        assert "(CUDA)" in ours["title"]
        want = {p["identifier"] for p in ref.get("parameters", [])}
        have = {p["identifier"] for p in ours.get("parameters", [])}
        assert want <= have, f"{name}: missing {sorted(want - have)}"
        for p in ours["parameters"]:
            assert p.get("title") and (p.get("type") or p["identifier"].endswith(".*")), p


def test_module_exports_mlt_register():
    out = subprocess.check_output(["nm", "-D", "--defined-only", mltenv.MODULE]).decode()
    assert " T mlt_register" in out


def dump(tmp_path, tag, cuda, profile, fmt, frames, args, rescale="bilinear", extra=None):
    prefix = str(tmp_path / tag)
    ex = {"MLT_DUMP_RESCALE": rescale}
    ex.update(extra or {})
    e = mltenv.env(str(tmp_path), cuda=True, override=cuda, extra=ex)
    r = subprocess.run([mltenv.DUMP, profile, fmt, prefix, ",".join(map(str, frames)), "--"] + args, env=e,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    res = {}
    for f in frames:
        img = np.fromfile(f"{prefix}.{f}.img", np.uint8)
        ap = f"{prefix}.{f}.alpha"
        res[f] = (img, np.fromfile(ap, np.uint8) if os.path.exists(ap) else None)
    return res, r.stderr


def test_module_fails_loudly_without_gpu(tmp_path):
    from mlt_b200 import abi
    if abi.load().mltcuda_device_count() > 0:
        pytest.skip("a GPU is visible")
    _, err = dump(tmp_path, "nogpu", True, "atsc_1080p_25", "yuv422", [3],
                  A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0/0:100%x100%:50%"])
    assert "no usable CUDA device" in err and "no CPU fallback" in err


# ---- GPU tier -------------------------------------------------------------------------------------

EXACT = [
    ("cfg1a_blit", "yuv422", [10], A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1"]),
    ("cfg1b_blend", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0/0:100%x100%:50%"]),
    ("cfg1c_subrect", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=25%/25%:50%x50%:50%"]),
    ("cfg1d_alpha_b", "yuv422", [10],
     A + ["-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
          "geometry=10%/10%:50%x50%:70%"]),
    # opaque colour B: the stock producer hands yuv422 WITHOUT an alpha plane (SSE2 case 1, not 3)
    ("comp_opaque_colour", "yuv422", [10],
     A + ["-track", "colour:0xc04020ff", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
          "geometry=10%/10%:50%x50%:70%"]),
    ("colour_checkerboard_rgba_out", "rgba", [3],
     ["colour:checkerboard", "out=249", "-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0",
      "b_track=1", "geometry=20%/20%:40%x40%:60%"]),
    ("comp_keyframed_odd", "yuv422", [0, 77, 249],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1",
                    "geometry=0=3/5:641x355:30%; 249=-101/-33:1501x901:90%"]),
    ("comp_luma_soft", "yuv422", [60],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0/0:100%x100%:60%",
                    f"luma={mltenv.LUMA_PGM}", "softness=0.3"]),
    ("comp_luma_invert", "yuv422", [60],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=10%/10%:80%x80%:40%",
                    f"luma={mltenv.LUMA_PGM}", "luma_invert=1"]),
    ("comp_xor_alpha", "yuv422", [10],
     A + ["-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
          "geometry=10%/10%:50%x50%:70%", "operator=xor"]),
    ("comp_halign_distort", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=5%/5%:60%x30%:75%",
                    "halign=centre", "valign=bottom"]),
    ("comp_crop", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=10%/10%:50%x50%:100%",
                    "crop=-100/-50:960x540"]),
    ("comp_fields", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0=0/0:50%x50%:50%; 249=50%/50%:50%x50%",
                    "progressive=0"]),
    ("cfg2_luma_wipe", "yuv422", [100],
     A + NOISE_B + ["-transition", f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "in=0", "out=249", "a_track=0", "b_track=1"]),
    ("luma_wipe_reverse_invert", "yuv422", [40],
     A + NOISE_B + ["-transition", f"luma:{mltenv.LUMA_PGM}", "softness=0.1", "reverse=1", "invert=1", "in=0", "out=249",
                    "a_track=0", "b_track=1"]),
    ("cfg2_dissolve", "yuv422", [100], A + NOISE_B + ["-transition", "luma", "in=0", "out=249", "a_track=0", "b_track=1"]),
    ("luma_wipe_rgba", "rgba", [100],
     A + NOISE_B + ["-transition", f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "in=0", "out=249", "a_track=0", "b_track=1"]),
    ("two_composites", "yuv422", [5],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-track",
                    "noise:", "out=249", "-transition", "composite", "a_track=0", "b_track=2",
                    "geometry=55%/55%:40%x40%:100%"]),
    ("imageconvert_rgba_out", "rgba", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0/0:100%x100%:50%"]),
    ("imageconvert_rgb_out", "rgb", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=25%/25%:50%x50%:50%"]),
    # stock wrapper filters that create the path's transitions BY NAME
    # (mlt_factory_transition(profile, "composite" / "luma")) pick up the CUDA services through the
    # create-request override -- core/filter_watermark.c, core/filter_luma.c
    ("wrapper_watermark", "yuv422", [10],
     A + ["-attach", "watermark:colour:0x3060c0a0", "composite.geometry=10%/10%:50%x50%:70%"]),
    ("wrapper_luma_filter", "yuv422", [13, 27],
     A + ["-attach", f"luma:{mltenv.LUMA_PGM}", "cycle=10", "duration=6"]),
    ("imageconvert_yuv420p_out", "yuv420p", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=25%/25%:50%x50%:50%"]),
    # ---- the loader's resize normaliser has to PAD (round-1 advisor finding): device-resident B whose aspect
    # differs from the request, no mirror:flip in between -> module/filter_rescale.c pads on the device
    ("resize_pad_colour_odd_rect", "yuv422", [0, 77],
     A + ["-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
          "geometry=0=3/5:641x355:30%; 249=-101/-33:1501x901:90%"]),
    ("resize_pad_noise_4x3_in_16x9", "yuv422", [10],
     A + ["-track", "noise:", "out=249", "meta.media.width=1440", "meta.media.height=1080", "-transition", "composite",
          "a_track=0", "b_track=1", "geometry=10%/10%:60%x60%:70%"]),
    ("resize_pad_colour_4x3_distort0", "yuv422", [10],
     A + ["-track", "colour:0x3060c0a0", "out=249", "meta.media.width=1440", "meta.media.height=1080", "-transition",
          "composite", "a_track=0", "b_track=1", "geometry=10%/10%:60%x60%:70%"]),
    # device rescale: B is a resident noise picture with a fixed native size that has to be scaled down
    ("rescale_noise_native_size", "yuv422", [10],
     A + ["-track", "noise:", "out=249", "meta.media.width=1920", "meta.media.height=1080", "-transition", "composite",
          "a_track=0", "b_track=1", "geometry=30%/30%:30%x30%:80%", "distort=1"]),
    # ---- composite properties the first round never exercised against the stock service (A5)
    ("comp_titles", "yuv422", [10],
     A + ["-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
          "geometry=10%/10%:50%x50%:70%", "titles=1"]),
    ("comp_crop_to_fill", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=10%/10%:30%x60%:80%", "crop_to_fill=1",
                    "halign=centre", "valign=middle"]),
    ("comp_fill", "yuv422", [10],
     A + ["-track", "noise:", "out=249", "meta.media.width=720", "meta.media.height=576", "-attach", "mirror:flip", "-transition",
          "composite", "a_track=0", "b_track=1", "geometry=5%/5%:90%x90%:80%", "fill=1"]),
    ("comp_distort", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=10%/10%:30%x60%:80%", "distort=1"]),
    ("comp_deinterlace", "yuv422", [10],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=10%/10%:50%x50%:80%", "deinterlace=1"]),
    ("comp_alpha_a_alpha_b", "yuv422", [10],
     ["colour:0x80402060", "out=249", "-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0",
      "b_track=1", "geometry=10%/10%:50%x50%:70%", "alpha_a=200", "alpha_b=90"]),
    ("comp_cycle_mirror_off", "yuv422", [0, 33, 77],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0=0/0:50%x50%:50%; 20=50%/50%:50%x50%",
                    "cycle=25", "mirror_off=1"]),
    ("comp_cycle_mirror", "yuv422", [33, 77],
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=0=0/0:50%x50%:50%; 20=50%/50%:50%x50%"]),
    ("comp_and_or", "yuv422", [10],
     ["colour:0x80402060", "out=249", "-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0",
      "b_track=1", "geometry=10%/10%:50%x50%:70%", "operator=and", "-track", "colour:0x10f0c040", "out=249", "-transition",
      "composite", "a_track=0", "b_track=2", "geometry=30%/30%:50%x50%:60%", "operator=or"]),
    # ---- the fused pass has to split: more ops than one k_tractor launch takes (MLTCUDA_TRACTOR_MAX_OPS = 8)
    # (rectangles are 16:9 with even heights: the reference's nearest scaler never writes the last row of an
    # odd-height output, filter_rescale.c:93 -- pool garbage no implementation can reproduce)
    ("ten_composites", "yuv422", [5], A + sum([["-track", f"colour:0x{(i * 37) % 256:02x}{(i * 91) % 256:02x}c0{255 - i * 20:02x}", "out=249",
                                                  "-transition", "composite", "a_track=0", f"b_track={i + 1}",
                                                  f"geometry={41 + i * 150}/{30 + i * 80}:{320 + 32 * i}x{180 + 18 * i}:{50 + i * 5}%"]
                                                 for i in range(10)], [])),
    # luma wipe fused with a composite on top (one k_tractor: luma, composite)
    ("luma_then_composite", "yuv422", [100],
     A + NOISE_B + ["-transition", f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "in=0", "out=249", "a_track=0", "b_track=1",
                    "-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=2",
                    "geometry=10%/10%:50%x50%:70%"]),
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,fmt,frames,args", EXACT, ids=[c[0] for c in EXACT])
def test_cuda_services_match_stock_services_exactly(tmp_path, name, fmt, frames, args):
    ref, _ = dump(tmp_path, "ref", False, "atsc_1080p_25", fmt, frames, args)
    got, err = dump(tmp_path, "cuda", True, "atsc_1080p_25", fmt, frames, args)
    assert "[cuda]" not in err.replace("[cuda] worker", ""), err[-1500:]
    for f in frames:
        assert ref[f][0].shape == got[f][0].shape
        bad = np.flatnonzero(ref[f][0] != got[f][0])
        assert bad.size == 0, f"frame {f}: {bad.size} differing bytes, first at {bad[:5]}"
        assert (ref[f][1] is None) == (got[f][1] is None)
        if ref[f][1] is not None:
            assert np.array_equal(ref[f][1], got[f][1])


AFFINE_B = ["-track", "noise:", "out=249", "meta.media.width=1920", "meta.media.height=1080", "-attach", "mirror:flip"]
CLOSE = [
    ("cfg3_affine_rgba", "rgba", [100], "bilinear",
     A + AFFINE_B + ["-transition", "affine", "rect=0=10%/10%:50%x50%:100%; 249=30%/30%:70%x70%:60%", "rotate_x=1", "in=0",
                     "out=249", "a_track=0", "b_track=1"]),
    ("cfg3_affine_yuv", "yuv422", [100], "bilinear",
     A + AFFINE_B + ["-transition", "affine", "rect=0=10%/10%:50%x50%:100%; 249=30%/30%:70%x70%:60%", "rotate_x=1", "in=0",
                     "out=249", "a_track=0", "b_track=1"]),
    ("affine_nearest_shear", "rgba", [30], "nearest",
     A + AFFINE_B + ["-transition", "affine", "rect=20%/10%:60%x60%:80%", "fix_shear_x=12", "fix_rotate_x=-20",
                     "a_track=0", "b_track=1"]),
    ("affine_bicubic_keyed", "rgba", [30], "bicubic",
     A + AFFINE_B + ["-transition", "affine", "rect=20%/10%:60%x60%:80%", "keyed=1", "rotate_x=0=0; 249=90", "ox=0=0; 249=300",
                     "a_track=0", "b_track=1"]),
    ("affine_colour_b_alpha", "rgba", [30], "bilinear",
     A + ["-track", "colour:0x3060c0a0", "out=249", "-transition", "affine", "rect=20%/20%:50%x50%:90%", "rotate_x=0.5",
          "b_alpha=1", "halign=centre", "valign=middle", "a_track=0", "b_track=1"]),
    # f3: plus/filter_affine.c instantiates the "affine" transition by name over a colour background
    ("wrapper_affine_filter", "rgba", [30], "bilinear",
     A + ["-attach", "affine", "transition.rect=10%/10%:60%x60%:80%", "transition.fix_rotate_x=15", "background=colour:0x20406080"]),
    # one B track read by two transitions: the composite's deferred op still holds the yuv422 handle when the affine
    # asks the same frame for rgba (a shared handle is never rewritten: the frame gets a converted copy)
    ("shared_b_track", "yuv422", [20], "bilinear",
     A + ["-track", "colour:0x3060c0a0", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
          "geometry=5%/5%:40%x40%:80%", "-transition", "affine", "a_track=0", "b_track=1", "rect=40%/40%:50%x50%:90%",
          "fix_rotate_x=10"]),
    ("shared_b_track_noise", "yuv422", [20], "bilinear",
     A + AFFINE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-transition",
                     "affine", "a_track=0", "b_track=1", "rect=40%/40%:50%x50%:90%", "fix_rotate_x=10", "-transition",
                     "composite", "a_track=0", "b_track=1", "geometry=60%/5%:30%x30%:60%"]),
    ("cfg4_1080p", "yuv422", [0, 37], "bilinear",
     A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-track",
                    "colour:0x3060c0a0", "out=599", "-transition", "affine", "a_track=0", "b_track=2",
                    "rect=0=20%/20%:50%x50%:90%; 599=30%/25%:60%x60%:70%", "rotate_x=0.5", "-track", "noise:", "out=599",
                    "-transition", "composite", "a_track=0", "b_track=3", "geometry=55%/55%:40%x40%:100%"]),
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,fmt,frames,rescale,args", CLOSE, ids=[c[0] for c in CLOSE])
def test_cuda_affine_matches_stock_within_1lsb(tmp_path, name, fmt, frames, rescale, args):
    ref, _ = dump(tmp_path, "ref", False, "atsc_1080p_25", fmt, frames, args, rescale)
    got, err = dump(tmp_path, "cuda", True, "atsc_1080p_25", fmt, frames, args, rescale)
    for f in frames:
        assert ref[f][0].shape == got[f][0].shape
        diff = np.abs(ref[f][0].astype(int) - got[f][0].astype(int))
        print(f"{name} frame {f}: {int((diff > 0).sum())} of {diff.size} bytes differ, max {int(diff.max())}")
        # spec: +-1 LSB (rgba).  The kernels follow the Release build's operation order, so in
        # practice frames are identical except where the double-precision x++ stepping of the
        # reference rounds differently from lower_x + j (a handful of pixels at most).
        assert diff.max() <= (1 if fmt == "rgba" else 2), f"frame {f}: max diff {diff.max()}"
        assert (diff > 0).mean() < 1e-4, f"frame {f}: {(diff > 0).mean():.6f} of bytes differ"
        if ref[f][1] is not None:
            assert got[f][1] is not None
            assert np.abs(ref[f][1].astype(int) - got[f][1].astype(int)).max() <= 1


@pytest.mark.gpu
def test_melt_renders_with_worker_threads(tmp_path):
    """melt ... -consumer null real_time=-4: four worker threads, one CUDA context each."""
    e = mltenv.env(str(tmp_path), cuda=True)
    args = A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=25%/25%:50%x50%:50%"]
    args = [a.replace("out=249", "out=59") for a in args]
    r = subprocess.run([mltenv.MELT, "-profile", "atsc_1080p_25"] + args + ["-consumer", "null", "real_time=-4", "-silent"],
                       env=e, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "no usable CUDA device" not in r.stderr and "failed" not in r.stderr


@pytest.mark.gpu
def test_device_frames_stay_resident_through_the_loader_normalisers(tmp_path):
    """melt attaches the loader's rescale/resize/... to every producer; the stock rescale forces yuv422 on
    what it pulls (filter_rescale.c:221-223).  The residency shim (module/filter_rescale.c) must keep the
    device producers' frames in HBM: no upload at all, one download per output frame."""
    n = 24
    args = ["noise:", "out=249", "-track", "noise:", "out=249", "-transition", "composite", "a_track=0", "b_track=1",
            "geometry=25%/25%:50%x50%:50%", "-track", "colour:0x3060c0a0", "out=249", "-transition", "affine", "a_track=0",
            "b_track=2", "rect=10%/10%:50%x50%:90%", "rotate_x=3"]
    stats, _ = melt_stats(tmp_path, "atsc_1080p_25", args, n)
    frames = stats.get("affine get_image (incl. inputs)")  # one per rendered frame
    assert frames and frames >= n, stats
    assert "upload" not in stats, stats
    assert stats.get("download + sync") == frames, stats
    assert stats.get("noise producer") == 2 * frames and stats.get("colour producer") == frames, stats


# ---- the fused tractor pass inside the module --------------------------------

CFG4 = A + NOISE_B + ["-transition", "composite", "a_track=0", "b_track=1", "geometry=5%/5%:40%x40%:80%", "-track",
                      "colour:0x3060c0a0", "out=599", "-transition", "affine", "a_track=0", "b_track=2",
                      "rect=0=20%/20%:50%x50%:90%; 599=30%/25%:60%x60%:70%", "rotate_x=0.5", "-track", "noise:", "out=599",
                      "-transition", "composite", "a_track=0", "b_track=3", "geometry=55%/55%:40%x40%:100%"]


def melt_stats(tmp_path, profile, args, frames, extra=None, workers=2):
    e = mltenv.env(str(tmp_path), cuda=True)
    e["MLT_CUDA_STATS"] = "1"
    e.update(extra or {})
    args = [a.replace("out=249", f"out={frames - 1}").replace("out=599", f"out={frames - 1}") for a in args]
    r = subprocess.run([mltenv.MELT, "-profile", profile] + args + ["-consumer", "null", f"real_time=-{workers}", "-silent"],
                       env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    stats = {}
    for ln in r.stderr.splitlines():
        if ln.startswith("[cuda stats]") and "calls" in ln:
            name, rest = ln[len("[cuda stats] "):].split("calls")
            stats[name.strip()] = int(rest.split()[0])
    assert stats, r.stderr[-2000:]
    return stats, r.stderr


@pytest.mark.gpu
def test_cfg4_through_melt_is_one_fused_launch_per_frame(tmp_path):
    """The drop-in defers composite / convert / affine / convert / composite onto A's op list and runs ONE
    k_tractor per frame when the consumer asks for the host image: no k_composite_rows / k_affine launch."""
    n = 24
    stats, err = melt_stats(tmp_path, "uhd_2160p_60", CFG4, n)
    frames = stats.get("affine get_image (incl. inputs)")  # one per rendered frame (melt re-renders the last one while it stops)
    assert frames and frames >= n, stats
    assert stats.get("fused pass (k_tractor launches)") == frames, stats
    assert stats.get("ops fused into k_tractor") == 5 * frames, stats
    for k in ("unfused composite launches", "unfused affine launches", "unfused luma launches"):
        assert k not in stats, stats
    assert "compositing launches per frame" in err


@pytest.mark.gpu
def test_fused_and_unfused_module_paths_are_byte_identical(tmp_path):
    fused, _ = dump(tmp_path, "fused", True, "atsc_1080p_25", "yuv422", [0, 37], CFG4)
    unfused, _ = dump(tmp_path, "unfused", True, "atsc_1080p_25", "yuv422", [0, 37], CFG4, extra={"MLT_CUDA_FUSE": "0"})
    for f in (0, 37):
        assert np.array_equal(fused[f][0], unfused[f][0])
        assert np.array_equal(fused[f][1], unfused[f][1])


def odd_profile(tmp_path, width=1004, height=562):
    """A profile whose width is not a multiple of 8: k_tractor refuses it (csrc/tractor.cu), the module must
    fall back to one launch per op."""
    p = tmp_path / "odd_profile"
    p.write_text(f"description=test {width}x{height}\nframe_rate_num=25\nframe_rate_den=1\nwidth={width}\nheight={height}\n"
                 "progressive=1\nsample_aspect_num=1\nsample_aspect_den=1\ndisplay_aspect_num=16\ndisplay_aspect_den=9\n"
                 "colorspace=709\n")
    return str(p)


@pytest.mark.gpu
def test_fused_pass_refusals_fall_back_to_per_op_launches(tmp_path):
    prof = odd_profile(tmp_path)
    # (a) width % 8 != 0: every op launches on its own, results still match the stock services
    ref, _ = dump(tmp_path, "ref", False, prof, "yuv422", [3], CFG4)
    got, err = dump(tmp_path, "cuda", True, prof, "yuv422", [3], CFG4)
    diff = np.abs(ref[3][0].astype(int) - got[3][0].astype(int))
    assert diff.max() <= 2 and (diff > 0).mean() < 1e-4
    stats, _ = melt_stats(tmp_path, prof, CFG4, 8)
    frames = stats["affine get_image (incl. inputs)"]
    assert "fused pass (k_tractor launches)" not in stats, stats
    assert stats.get("unfused composite launches") == 2 * frames, stats
    assert stats.get("unfused affine launches") == frames, stats
    # (b) rgba64 working format: the affine launches K4, the composites around it stay deferred
    a64 = ["abnormal:colour:0x80c04020", "out=249", "-track", "abnormal:colour:0x3060c0a0", "out=249", "-transition", "affine",
           "rect=20%/20%:50%x50%:90%", "rotate_x=0.5", "a_track=0", "b_track=1"]
    ref, _ = dump(tmp_path, "ref64", False, "atsc_1080p_25", "rgba64", [30], a64)
    got, _ = dump(tmp_path, "cuda64", True, "atsc_1080p_25", "rgba64", [30], a64)
    r16, g16 = ref[30][0].view(np.uint16).astype(int), got[30][0].view(np.uint16).astype(int)
    assert r16.shape == g16.shape and np.abs(r16 - g16).max() <= 257


# ---- translucent (float) luma paths pinned to the compiled reference -------------------

TRANSLUCENT_B = ["-track", "colour:0x3060c0a0", "out=249"]
LUMA_FLOAT = [
    ("wipe_yuv422_translucent", "yuv422", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2"]),
    ("wipe_yuv422_translucent_over", "yuv422", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "alpha_over=1"]),
    ("dissolve_yuv422_translucent", "yuv422", ["luma"]),
    ("dissolve_yuv422_translucent_over", "yuv422", ["luma", "alpha_over=1"]),
    ("wipe_rgba_translucent", "rgba", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2"]),
    ("wipe_rgba_translucent_over", "rgba", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "alpha_over=1"]),
    ("dissolve_rgba_translucent_over", "rgba", ["luma", "alpha_over=1"]),
    ("wipe_rgba64_translucent", "rgba64", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2"]),
    ("wipe_rgba64_translucent_over", "rgba64", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "alpha_over=1"]),
    ("dissolve_rgba64_translucent_over", "rgba64", ["luma", "alpha_over=1"]),
    ("wipe_yuv422_fix_background", "yuv422", [f"luma:{mltenv.LUMA_PGM}", "softness=0.2", "fix_background_alpha=1"]),
    ("dissolve_yuv422_fix_background", "yuv422", ["luma", "fix_background_alpha=1"]),
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,fmt,trans", LUMA_FLOAT, ids=[c[0] for c in LUMA_FLOAT])
def test_translucent_luma_matches_stock_within_1lsb(tmp_path, name, fmt, trans):
    """B = colour:0x3060c0a0 (alpha 0xa0): is_opaque() fails, the reference takes its float paths
    (transition_luma.c:63-129,192-292,417-451,475-723).  The colour tracks are rgba64-native for the rgba64 cases."""
    # rgba64: the loader's rescale normaliser forces yuv422 and this build has no yuv422 -> rgba64 converter
    # (swscale absent), so the stock services only reach rgba64 through "abnormal:" (loader without normalisers)
    a = ["abnormal:colour:0x80c04020", "out=249"] if fmt == "rgba64" else A
    b = ["-track", "abnormal:colour:0x3060c0a0", "out=249"] if fmt == "rgba64" else TRANSLUCENT_B
    args = a + b + ["-transition"] + trans + ["in=0", "out=249", "a_track=0", "b_track=1"]
    ref, _ = dump(tmp_path, "ref", False, "atsc_1080p_25", fmt, [60, 130], args)
    got, err = dump(tmp_path, "cuda", True, "atsc_1080p_25", fmt, [60, 130], args)
    for f in (60, 130):
        if fmt == "rgba64":
            r, g = ref[f][0].view(np.uint16).astype(int), got[f][0].view(np.uint16).astype(int)
            tol = 257  # one 8-bit LSB
        else:
            r, g = ref[f][0].astype(int), got[f][0].astype(int)
            tol = 1
        assert r.shape == g.shape
        d = np.abs(r - g)
        print(f"{name} frame {f}: {int((d > 0).sum())} of {d.size} differ, max {int(d.max())}")
        assert d.max() <= tol, f"frame {f}: max diff {d.max()}"
        assert (ref[f][1] is None) == (got[f][1] is None)
        if ref[f][1] is not None:
            assert np.abs(ref[f][1].astype(int) - got[f][1].astype(int)).max() <= 1


# ---- affine properties the first round never exercised against the stock service (F4) -----------------------

AFFINE_PROPS = [
    ("scale_xy", ["rect=20%/10%:60%x60%:80%", "scale_x=1.5", "scale_y=0.75"]),
    ("invert_scale", ["rect=20%/10%:60%x60%:80%", "scale_x=1.5", "scale_y=0.75", "invert_scale=1"]),
    ("mirror", ["rect=0=10%/10%:50%x50%:100%; 100=30%/30%:70%x70%:60%", "mirror=1", "in=0", "out=199"]),
    ("rotate_y_z", ["rect=20%/10%:60%x60%:80%", "fix_rotate_y=25", "fix_rotate_z=-15"]),
    ("b_scaled", ["rect=20%/10%:40%x40%:80%", "b_scaled=1"]),
    ("fill0", ["rect=20%/10%:60%x40%:80%", "fill=0"]),
    ("distort", ["rect=20%/10%:60%x30%:80%", "distort=1"]),
    ("scale_flag", ["rect=20%/10%:60%x60%:80%", "fix_rotate_x=30", "scale=1"]),
    ("cycle", ["rect=0=10%/10%:50%x50%:100%; 20=30%/30%:70%x70%:60%", "cycle=25"]),
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,props", AFFINE_PROPS, ids=[c[0] for c in AFFINE_PROPS])
def test_affine_properties_match_stock(tmp_path, name, props):
    args = A + AFFINE_B + ["-transition", "affine"] + props + ["a_track=0", "b_track=1"]
    frames = [30, 160]
    ref, _ = dump(tmp_path, "ref", False, "atsc_1080p_25", "rgba", frames, args)
    got, _ = dump(tmp_path, "cuda", True, "atsc_1080p_25", "rgba", frames, args)
    for f in frames:
        d = np.abs(ref[f][0].astype(int) - got[f][0].astype(int))
        print(f"{name} frame {f}: {int((d > 0).sum())} of {d.size} differ, max {int(d.max())}")
        assert d.max() <= 1 and (d > 0).mean() < 1e-4


@pytest.mark.gpu
def test_foreign_rescaler_keeps_its_own_interpolation(tmp_path):
    """swscale / gtkrescale are core `rescale` with the scaler function swapped, and the loader prefers them.  The
    device nearest-neighbour scaler (K6) must not answer for them: mlt_dump's stand-in (MLT_DUMP_FOREIGN_RESCALER,
    a scaler no other code path can imitate) has to produce the same picture with and without the module, under a
    yuv422 composite and under an rgba affine."""
    ex = {"MLT_DUMP_FOREIGN_RESCALER": "1"}
    comp = A + ["-track", "noise:", "out=249", "meta.media.width=1920", "meta.media.height=1080", "-transition", "composite",
                "a_track=0", "b_track=1", "geometry=30%/30%:30%x30%:80%", "distort=1"]
    plain, _ = dump(tmp_path, "plain", False, "atsc_1080p_25", "yuv422", [10], comp)
    ref, _ = dump(tmp_path, "ref", False, "atsc_1080p_25", "yuv422", [10], comp, extra=ex)
    got, err = dump(tmp_path, "cuda", True, "atsc_1080p_25", "yuv422", [10], comp, extra=ex)
    assert not np.array_equal(plain[10][0], ref[10][0]), "the stand-in scaler was not picked up by the loader"
    assert np.array_equal(ref[10][0], got[10][0]), f"{(ref[10][0] != got[10][0]).sum()} bytes differ"
    aff = A + AFFINE_B + ["-transition", "affine", "rect=20%/10%:40%x40%:80%", "b_scaled=1", "a_track=0", "b_track=1"]
    ref, _ = dump(tmp_path, "ref", False, "atsc_1080p_25", "rgba", [30], aff, extra=ex)
    got, _ = dump(tmp_path, "cuda", True, "atsc_1080p_25", "rgba", [30], aff, extra=ex)
    d = np.abs(ref[30][0].astype(int) - got[30][0].astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-4, f"max {d.max()}, {(d > 0).sum()} differ"


@pytest.mark.gpu
def test_cuda_null_consumer_takes_device_frames(tmp_path):
    """-consumer cuda.null asks for mlt_image_private; frames stay in HBM (no download), the
    deferred list still runs once per frame."""
    e = mltenv.env(str(tmp_path), cuda=True)
    e["MLT_CUDA_STATS"] = "1"
    n = 24
    args = [a.replace("out=249", f"out={n - 1}").replace("out=599", f"out={n - 1}") for a in CFG4 if a != "mirror:flip" and a != "-attach"]
    r = subprocess.run([mltenv.MELT, "-profile", "uhd_2160p_60"] + args + ["-consumer", "cuda.null", "real_time=-2", "terminate_on_pause=1",
                                                                            "-silent"], env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    stats = {}
    for ln in r.stderr.splitlines():
        if ln.startswith("[cuda stats]") and "calls" in ln:
            name, rest = ln[len("[cuda stats] "):].split("calls")
            stats[name.strip()] = int(rest.split()[0])
    frames = stats.get("affine get_image (incl. inputs)")
    assert frames and frames >= n, (stats, r.stderr[-1500:])
    assert "download + sync" not in stats and "upload" not in stats, stats
    # (frames melt pulls again while it stops are dropped before their hand-off)
    assert n <= stats.get("fused pass (k_tractor launches)") <= frames, stats


@pytest.mark.gpu
def test_cuda_null_hands_off_only_to_itself(tmp_path):
    """A stock CPU filter planted above the last CUDA transition (greyscale on track 0) asks the rendered frame for
    yuv422 before the consumer does.  It must get pixels: cuda.null may take the device handle only when its own
    request is provably the first host-format one (consumer_null.c: consumer_is_sole_requester), so here every
    frame is downloaded for the filter and nothing is handed off."""
    e = mltenv.env(str(tmp_path), cuda=True)
    e["MLT_CUDA_STATS"] = "1"
    n = 12
    args = ["noise:", f"out={n - 1}", "-track", "noise:", f"out={n - 1}", "-transition", "composite", "a_track=0", "b_track=1",
            "geometry=25%/25%:50%x50%:50%", "-filter", "greyscale", "track=0"]
    r = subprocess.run([mltenv.MELT, "-profile", "atsc_1080p_25"] + args + ["-consumer", "cuda.null", "real_time=-2",
                                                                            "terminate_on_pause=1", "-silent"],
                       env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    stats = {}
    for ln in r.stderr.splitlines():
        if ln.startswith("[cuda stats]") and "calls" in ln:
            name, rest = ln[len("[cuda stats] "):].split("calls")
            stats[name.strip()] = int(rest.split()[0])
    frames = stats.get("composite get_image (incl. inputs)")
    assert frames and frames >= n, (stats, r.stderr[-1500:])
    assert stats.get("download + sync") == frames, stats
