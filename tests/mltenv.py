"""Environment for running the reference's melt / mlt_dump (oracle/_ref) with or without the CUDA
module (mlt_b200/lib/mlt/libmltcuda.so) loaded.  Test infrastructure.

The combined repository directory holds symlinks to the stock modules plus libmltcuda.so, and the
combined data directory holds the stock data plus cuda/*.yml, so nothing under oracle/_ref is
modified.  MLT_CUDA_OVERRIDE=1 (default) makes `composite`, `luma`, `affine`, `imageconvert`
resolve to the CUDA services; =0 leaves the stock services and only adds the cuda.* aliases.
"""
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
MODULE = os.path.join(ROOT, "mlt_b200", "lib", "mlt", "libmltcuda.so")
RT_DIR = os.path.join(ROOT, "mlt_b200", "lib")
DUMP = os.path.join(REF, "bin", "mlt_dump")
MELT = os.path.join(REF, "bin", "melt")
LUMA_PGM = os.path.join(REF, "share", "mlt", "demo", "luma1.pgm")


def have_reference():
    return os.path.exists(os.path.join(REF, "lib", "mlt", "libmltcore.so")) and os.path.exists(DUMP)


def have_module():
    return os.path.exists(MODULE)


def combined_dirs(base):
    """Create <base>/repo and <base>/data; returns (repo, data)."""
    repo, data = os.path.join(base, "repo"), os.path.join(base, "data")
    os.makedirs(repo, exist_ok=True)
    for f in os.listdir(os.path.join(REF, "lib", "mlt")):
        dst = os.path.join(repo, f)
        if not os.path.lexists(dst):
            os.symlink(os.path.join(REF, "lib", "mlt", f), dst)
    dst = os.path.join(repo, "libmltcuda.so")
    if not os.path.lexists(dst):
        os.symlink(MODULE, dst)
    if not os.path.isdir(data):
        os.makedirs(data)
        for f in os.listdir(os.path.join(REF, "share", "mlt")):
            os.symlink(os.path.join(REF, "share", "mlt", f), os.path.join(data, f))
        shutil.copytree(os.path.join(ROOT, "mlt_b200", "lib", "share", "cuda"), os.path.join(data, "cuda"))
    return repo, data


def env(base=None, cuda=False, override=True, extra=None):
    e = dict(os.environ)
    if cuda:
        repo, data = combined_dirs(base)
    else:
        repo, data = os.path.join(REF, "lib", "mlt"), os.path.join(REF, "share", "mlt")
    e["MLT_REPOSITORY"] = repo
    e["MLT_DATA"] = data
    e["MLT_PROFILES_PATH"] = os.path.join(REF, "share", "mlt", "profiles")
    e["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(REF, "lib"), RT_DIR, e.get("LD_LIBRARY_PATH", "")])
    e["MLT_CUDA_OVERRIDE"] = "1" if override else "0"
    if extra:
        e.update(extra)
    return e
