"""mlt_b200 -- B200-native (sm_100a) implementation of MLT's per-frame compositing hot path.

Product artefacts (built in-tree by `make -C mlt_b200/csrc` and `make -C mlt_b200/module`):
  mlt_b200/lib/libmltcuda_rt.so   hand-written CUDA kernels behind the C-ABI of include/mltcuda.h
  mlt_b200/lib/mlt/libmltcuda.so  the MLT module (C) registering composite/luma/affine/imageconvert
The Python in this package is only the ctypes plumbing tests/ and bench.py use.
"""
from . import abi  # noqa: F401
from .device import Context  # noqa: F401
