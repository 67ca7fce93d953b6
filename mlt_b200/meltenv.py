"""Environment for running an MLT host (melt, or any libmlt-7 application) with the CUDA module loaded.

MLT finds modules in $MLT_REPOSITORY and their metadata in $MLT_DATA (src/framework/mlt_factory.c:116-175).
The module is dropped into an existing install by pointing both at directories that hold the stock
entries (symlinks) plus mlt_b200/lib/mlt/libmltcuda.so and mlt_b200/lib/share/cuda/*.yml -- nothing of
the MLT install is modified.  `prefix` is that install; here the only one available is
the reference build under oracle/_ref (any MLT 7 prefix works the same way).
"""
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_PREFIX = os.path.join(ROOT, "oracle", "_ref")
MODULE = os.path.join(ROOT, "mlt_b200", "lib", "mlt", "libmltcuda.so")
RT_DIR = os.path.join(ROOT, "mlt_b200", "lib")


def melt_binary(prefix=DEFAULT_PREFIX):
    return os.path.join(prefix, "bin", "melt")


def have_mlt(prefix=DEFAULT_PREFIX):
    return os.path.exists(melt_binary(prefix)) and os.path.exists(os.path.join(prefix, "lib", "mlt", "libmltcore.so"))


def have_module():
    return os.path.exists(MODULE)


def combined_dirs(base, prefix=DEFAULT_PREFIX):
    """Create <base>/repo (stock modules + libmltcuda.so) and <base>/data (stock data + cuda/); returns both."""
    repo, data = os.path.join(base, "repo"), os.path.join(base, "data")
    os.makedirs(repo, exist_ok=True)
    for f in os.listdir(os.path.join(prefix, "lib", "mlt")):
        dst = os.path.join(repo, f)
        if not os.path.lexists(dst):
            os.symlink(os.path.join(prefix, "lib", "mlt", f), dst)
    dst = os.path.join(repo, "libmltcuda.so")
    if not os.path.lexists(dst):
        os.symlink(MODULE, dst)
    if not os.path.isdir(data):
        os.makedirs(data)
        for f in os.listdir(os.path.join(prefix, "share", "mlt")):
            os.symlink(os.path.join(prefix, "share", "mlt", f), os.path.join(data, f))
        shutil.copytree(os.path.join(ROOT, "mlt_b200", "lib", "share", "cuda"), os.path.join(data, "cuda"))
    return repo, data


def env(base=None, cuda=False, override=True, extra=None, prefix=DEFAULT_PREFIX):
    """Process environment for the MLT host.  cuda=False: the stock install alone.  cuda=True: the module is
    loaded; override=True makes `composite`, `luma`, `affine`, `imageconvert` resolve to it, False leaves
    the stock services and only adds the cuda.* aliases."""
    e = dict(os.environ)
    if cuda:
        repo, data = combined_dirs(base, prefix)
    else:
        repo, data = os.path.join(prefix, "lib", "mlt"), os.path.join(prefix, "share", "mlt")
    e["MLT_REPOSITORY"] = repo
    e["MLT_DATA"] = data
    e["MLT_PROFILES_PATH"] = os.path.join(prefix, "share", "mlt", "profiles")
    e["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(prefix, "lib"), RT_DIR, e.get("LD_LIBRARY_PATH", "")])
    e["MLT_CUDA_OVERRIDE"] = "1" if override else "0"
    if extra:
        e.update(extra)
    return e
