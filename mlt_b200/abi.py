"""ctypes binding of include/mltcuda.h (the C-ABI of mlt_b200/lib/libmltcuda_rt.so).

This is plumbing for tests/ and bench.py only: the product is the C-ABI library and the MLT module
(mlt_b200/module -> libmltcuda.so) on top of it.  There is no CPU fallback: if the CUDA library is
missing, or no GPU is visible when a context is created, this raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MLTCUDA_RT_LIB") or os.path.join(HERE, "lib", "libmltcuda_rt.so")  # (override: A/B builds of tools/ab.sh)

FMT_RGB, FMT_RGBA, FMT_YUV422, FMT_YUV420P, FMT_RGBA64 = 1, 2, 3, 4, 10
OP_OVER, OP_OR, OP_AND, OP_XOR = 0, 1, 2, 3
INTERP_NN, INTERP_BL, INTERP_BC = 0, 1, 2
TOP_COMPOSITE, TOP_AFFINE, TOP_LUMA, TOP_CONVERT = 1, 2, 3, 4
E_BADARG, E_UNSUPPORTED = 10001, 10002
TRACTOR_MAX_OPS = 8

vp, sz = C.c_void_p, C.c_size_t


class ConvertDesc(C.Structure):
    _fields_ = [("src_format", C.c_int), ("dst_format", C.c_int), ("width", C.c_int), ("height", C.c_int),
                ("src", vp), ("dst", vp), ("src_alpha", vp), ("dst_alpha", vp),
                ("src_stride", C.c_int), ("dst_stride", C.c_int), ("batch", C.c_int),
                ("src_bstride", sz), ("dst_bstride", sz), ("src_alpha_bstride", sz), ("dst_alpha_bstride", sz)]


class CompositeDesc(C.Structure):
    _fields_ = [("dest", vp), ("src", vp), ("alpha_b", vp), ("alpha_a", vp), ("luma", vp),
                ("width", C.c_int), ("rows", C.c_int), ("stride_dest", C.c_int), ("stride_src", C.c_int),
                ("stride_alpha_a", C.c_int), ("stride_alpha_b", C.c_int),
                ("weight", C.c_int), ("softness", C.c_int), ("step", C.c_uint32),
                ("op", C.c_int), ("scalar_only", C.c_int), ("batch", C.c_int),
                ("dest_bstride", sz), ("src_bstride", sz), ("alpha_a_bstride", sz),
                ("alpha_b_bstride", sz), ("luma_bstride", sz)]


class CompositeGeometry(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("w", C.c_double), ("h", C.c_double), ("o", C.c_double),
                ("nw", C.c_int), ("nh", C.c_int), ("sw", C.c_int), ("sh", C.c_int),
                ("x_src", C.c_int), ("y_src", C.c_int)]


class LumaDesc(C.Structure):
    _fields_ = [("format", C.c_int), ("dest", vp), ("src", vp), ("alpha_dest", vp), ("alpha_src", vp),
                ("map", vp), ("luma_width", C.c_int), ("luma_height", C.c_int),
                ("width", C.c_int), ("height", C.c_int),
                ("width_dest", C.c_int), ("height_dest", C.c_int), ("width_src", C.c_int), ("height_src", C.c_int),
                ("pos", C.c_float), ("frame_delta", C.c_float), ("softness", C.c_float),
                ("field_order", C.c_int), ("invert", C.c_int), ("alpha_over", C.c_int),
                ("translucent", C.c_int), ("batch", C.c_int),
                ("dest_bstride", sz), ("src_bstride", sz), ("alpha_dest_bstride", sz), ("alpha_src_bstride", sz)]


class AffineDesc(C.Structure):
    _fields_ = [("format", C.c_int), ("a_image", vp), ("b_image", vp),
                ("a_width", C.c_int), ("a_height", C.c_int), ("b_width", C.c_int), ("b_height", C.c_int),
                ("matrix", (C.c_double * 3) * 3),
                ("lower_x", C.c_double), ("lower_y", C.c_double), ("dz", C.c_double), ("mix", C.c_double),
                ("x_offset", C.c_double), ("y_offset", C.c_double),
                ("minima", C.c_double), ("xmax", C.c_double), ("ymax", C.c_double),
                ("b_alpha", C.c_int), ("interp", C.c_int), ("batch", C.c_int),
                ("a_bstride", sz), ("b_bstride", sz)]


class _OpUnion(C.Union):
    _fields_ = [("composite", CompositeDesc), ("affine", AffineDesc), ("luma", LumaDesc)]


class TractorOp(C.Structure):
    _fields_ = [("kind", C.c_int), ("to_format", C.c_int), ("u", _OpUnion)]


class TractorDesc(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("in_format", C.c_int), ("out_format", C.c_int),
                ("in_", vp), ("in_alpha", vp), ("out", vp), ("out_alpha", vp),
                ("n_ops", C.c_int), ("ops", C.POINTER(TractorOp)), ("batch", C.c_int),
                ("in_bstride", sz), ("in_alpha_bstride", sz), ("out_bstride", sz), ("out_alpha_bstride", sz)]


# every symbol include/mltcuda.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "mltcuda_abi_version": (C.c_int, []),
    "mltcuda_device_count": (C.c_int, []),
    "mltcuda_last_error": (C.c_char_p, []),
    "mltcuda_ctx_create": (vp, [C.c_int]),
    "mltcuda_ctx_destroy": (None, [vp]),
    "mltcuda_ctx_device": (C.c_int, [vp]),
    "mltcuda_ctx_stream": (vp, [vp]),
    "mltcuda_ctx_sync": (C.c_int, [vp]),
    "mltcuda_ctx_launches": (C.c_uint64, [vp]),
    "mltcuda_malloc": (vp, [vp, sz]),
    "mltcuda_free": (None, [vp, vp]),
    "mltcuda_host_alloc": (vp, [sz]),
    "mltcuda_host_free": (None, [vp]),
    "mltcuda_host_register": (C.c_int, [vp, sz]),
    "mltcuda_host_unregister": (C.c_int, [vp]),
    "mltcuda_upload": (C.c_int, [vp, vp, vp, sz]),
    "mltcuda_download": (C.c_int, [vp, vp, vp, sz]),
    "mltcuda_copy": (C.c_int, [vp, vp, vp, sz]),
    "mltcuda_memset": (C.c_int, [vp, vp, C.c_int, sz]),
    "mltcuda_graph_begin": (C.c_int, [vp]),
    "mltcuda_graph_end": (vp, [vp]),
    "mltcuda_graph_launch": (C.c_int, [vp, vp]),
    "mltcuda_graph_kernel_nodes": (C.c_uint64, [vp]),
    "mltcuda_graph_destroy": (None, [vp]),
    "mltcuda_plane_opaque_query": (C.c_int, [vp, vp, sz]),
    "mltcuda_plane_opaque_result": (C.c_int, [vp]),
    "mltcuda_plane_opaque_query_into": (C.c_int, [vp, vp, sz, vp]),
    "mltcuda_event_create": (vp, []),
    "mltcuda_event_destroy": (None, [vp]),
    "mltcuda_event_record": (C.c_int, [vp, vp]),
    "mltcuda_event_sync": (C.c_int, [vp]),
    "mltcuda_ctx_wait_event": (C.c_int, [vp, vp]),
    "mltcuda_ctx_wait_ctx": (C.c_int, [vp, vp]),
    "mltcuda_copy_peer": (C.c_int, [vp, vp, vp, vp, sz]),
    "mltcuda_upload_staged": (C.c_int, [vp, vp, vp, C.c_size_t]),
    "mltcuda_host_pool_alloc": (vp, [C.c_size_t]),
    "mltcuda_host_pool_release": (None, [vp]),
    "mltcuda_event_elapsed_ms": (C.c_float, [vp, vp]),
    "mltcuda_convert": (C.c_int, [vp, C.POINTER(ConvertDesc)]),
    "mltcuda_composite": (C.c_int, [vp, C.POINTER(CompositeDesc)]),
    "mltcuda_composite_setup": (C.c_int, [C.POINTER(CompositeDesc), vp, C.c_int, C.c_int, vp, C.c_int, C.c_int,
                                          vp, vp, C.POINTER(CompositeGeometry), C.c_int, vp, C.c_double]),
    "mltcuda_luma": (C.c_int, [vp, C.POINTER(LumaDesc)]),
    "mltcuda_scale_luma": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int]),
    "mltcuda_affine": (C.c_int, [vp, C.POINTER(AffineDesc)]),
    "mltcuda_tractor": (C.c_int, [vp, C.POINTER(TractorDesc)]),
    "mltcuda_noise": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, sz]),
    "mltcuda_noise_static": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, sz]),
    "mltcuda_colour": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_uint8, C.c_uint8, C.c_uint8, C.c_uint8]),
    "mltcuda_rescale_yuv422": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int, sz, sz]),
    "mltcuda_rescale_alpha": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, C.c_int, C.c_int]),
    "mltcuda_mirror_flip": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int]),
    "mltcuda_resize_pad": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int]),
}

_lib = None


def load():
    """Load libmltcuda_rt.so and type every entry point.  Raises if the library is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C mlt_b200/csrc` (or __graft_entry__.build()); "
                "there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


class MltCudaError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != 0:
        raise MltCudaError(f"{what} failed rc={rc}: {load().mltcuda_last_error().decode()}")
