"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports exactly the entry
points include/mltcuda.h declares (no compute calls: there is no GPU here)."""
import os
import re

import pytest

from mlt_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mltcuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mltcuda_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported_and_typed():
    lib = abi.load()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/mltcuda.h but not exported"
        assert n in abi.SIGNATURES, f"{n} has no ctypes signature"
    assert sorted(abi.SIGNATURES) == names


def test_abi_version_and_no_device_error():
    lib = abi.load()
    assert lib.mltcuda_abi_version() == 2
    if lib.mltcuda_device_count() == 0:
        # the product path fails loudly without a GPU: no CPU fallback
        assert not lib.mltcuda_ctx_create(0)
        assert b"no CUDA device" in lib.mltcuda_last_error()


def test_struct_sizes_match_header():
    """sizeof() of every descriptor as the C compiler sees it == the ctypes mirror."""
    import ctypes as C
    import subprocess
    import tempfile
    src = r'''
#include <stdio.h>
#include "mltcuda.h"
int main(void){printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(mltcuda_convert_desc), sizeof(mltcuda_composite_desc),
 sizeof(mltcuda_composite_geometry), sizeof(mltcuda_luma_desc), sizeof(mltcuda_affine_desc), sizeof(mltcuda_tractor_op),
 sizeof(mltcuda_tractor_desc));return 0;}'''
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(td, "s"), os.path.join(td, "s.c")])
        out = subprocess.check_output([os.path.join(td, "s")]).split()
    sizes = [int(x) for x in out]
    mirror = [abi.ConvertDesc, abi.CompositeDesc, abi.CompositeGeometry, abi.LumaDesc, abi.AffineDesc, abi.TractorOp,
              abi.TractorDesc]
    assert sizes == [C.sizeof(m) for m in mirror]


def test_no_contracted_pair_fma():
    """ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under --fmad=false; the
    samplers must keep the reference's two roundings, so no kernel may contain FFMA2 it did not ask for."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mlt_b200", "lib", "libmltcuda_rt.so")
    sass = subprocess.run([cuobjdump, "-sass", so], capture_output=True, text=True, check=True).stdout
    assert "FADD2" in sass and "FMUL2" in sass, "packed fp32 pair ops missing: wrong build?"
    # per kernel: FFMA2 is written on purpose only in k_luma_rgba / k_luma_yuv422 (the packed correctly rounded divisions,
    # mc_fdiv_const2 / mc_fdiv_fast_rn2, fused in the scalar code too); anywhere else it is a contraction
    fn, bad = "", set()
    for ln in sass.splitlines():
        if "Function :" in ln:
            fn = ln.split("Function :")[1].strip()
        elif "FFMA2" in ln and "k_luma_" not in fn:
            bad.add(fn)
    assert not bad, f"a packed multiply-add pair was contracted (results would differ from the reference) in: {sorted(bad)}"
