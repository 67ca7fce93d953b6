"""Thin numpy <-> device helpers over the C-ABI (tests/ and bench.py plumbing)."""
import ctypes as C

import numpy as np

from . import abi


class DevBuf:
    def __init__(self, ctx, nbytes):
        self.ctx, self.nbytes = ctx, int(nbytes)
        self.ptr = abi.load().mltcuda_malloc(ctx.h, self.nbytes)
        if not self.ptr:
            raise abi.MltCudaError("mltcuda_malloc: " + abi.load().mltcuda_last_error().decode())

    def free(self):
        if self.ptr:
            abi.load().mltcuda_free(self.ctx.h, self.ptr)
            self.ptr = None

    def at(self, byte_offset):
        return self.ptr + int(byte_offset)


class Context:
    """One mltcuda_ctx (stream + pool) on one GPU.  Raises if no GPU / library."""

    def __init__(self, device=0):
        self.lib = abi.load()
        self.h = self.lib.mltcuda_ctx_create(device)
        if not self.h:
            raise abi.MltCudaError("mltcuda_ctx_create: " + self.lib.mltcuda_last_error().decode())
        self._bufs = []

    def close(self):
        for b in self._bufs:
            b.free()
        self._bufs = []
        if self.h:
            self.lib.mltcuda_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def alloc(self, nbytes):
        b = DevBuf(self, nbytes)
        self._bufs.append(b)
        return b

    def put(self, arr):
        """Upload a numpy array; returns DevBuf."""
        arr = np.ascontiguousarray(arr)
        b = self.alloc(max(arr.nbytes, 1))
        if arr.nbytes:
            abi.check(self.lib.mltcuda_upload(self.h, b.ptr, arr.ctypes.data, arr.nbytes), "upload")
            self.sync()  # pageable source: keep it simple and safe
        return b

    def get(self, buf, shape, dtype=np.uint8, offset=0):
        out = np.empty(shape, dtype=dtype)
        if out.nbytes:
            abi.check(self.lib.mltcuda_download(self.h, out.ctypes.data, buf.at(offset), out.nbytes), "download")
            self.sync()
        return out

    def sync(self):
        abi.check(self.lib.mltcuda_ctx_sync(self.h), "sync")

    def launches(self):
        return int(self.lib.mltcuda_ctx_launches(self.h))

    def call(self, name, desc):
        abi.check(getattr(self.lib, name)(self.h, C.byref(desc)), name)
