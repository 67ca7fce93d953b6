import sys, os, tempfile, pathlib
sys.path.insert(0, "tests")
import numpy as np
import test_module as tm
A = tm.A
def comps(n):
    return A + sum([["-track", f"colour:0x{(i * 37) % 256:02x}{(i * 91) % 256:02x}c0{255 - i * 20:02x}", "out=249",
                     "-transition", "composite", "a_track=0", f"b_track={i + 1}",
                     f"geometry={3 + i * 8}%/{5 + i * 7}%:{20 + i}%x{18 + i}%:{50 + i * 5}%"] for i in range(n)], [])
for n in (2, 3, 4, 8, 9, 10):
    tmp = pathlib.Path(tempfile.mkdtemp())
    ref, _ = tm.dump(tmp, "ref", False, "atsc_1080p_25", "yuv422", [5], comps(n))
    for fuse in ("1", "0"):
        got, err = tm.dump(tmp, "cuda" + fuse, True, "atsc_1080p_25", "yuv422", [5], comps(n), extra={"MLT_CUDA_FUSE": fuse})
        bad = np.flatnonzero(ref[5][0] != got[5][0])
        rows = sorted(set((bad // 3840).tolist()))
        xs = sorted(set(((bad % 3840) // 2).tolist()))
        print(n, "fuse", fuse, "bad", bad.size, "rows", rows[:3], rows[-3:], "x", xs[:4], xs[-4:], "alpha", (ref[5][1] is None, got[5][1] is None),
              "" if ref[5][1] is None or got[5][1] is None else int((ref[5][1] != got[5][1]).sum()))
        if bad.size:
            d = ref[5][0].astype(int)[bad] - got[5][0].astype(int)[bad]
            print("   diffs", np.unique(d)[:10], err[-300:])
