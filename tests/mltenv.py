"""Environment for running the reference's melt / mlt_dump (oracle/_ref) with or without the CUDA
module (mlt_b200/lib/mlt/libmltcuda.so) loaded.  Test infrastructure; the mechanics live in
mlt_b200/meltenv.py (the same recipe INTEGRATION.md gives for any MLT 7 install).
"""
import os

from mlt_b200 import meltenv

ROOT = meltenv.ROOT
REF = meltenv.DEFAULT_PREFIX
MODULE = meltenv.MODULE
RT_DIR = meltenv.RT_DIR
DUMP = os.path.join(REF, "bin", "mlt_dump")
MELT = meltenv.melt_binary()
LUMA_PGM = os.path.join(REF, "share", "mlt", "demo", "luma1.pgm")


def have_reference():
    return meltenv.have_mlt() and os.path.exists(DUMP)


have_module = meltenv.have_module
combined_dirs = meltenv.combined_dirs
env = meltenv.env
