/*
 * mltcuda.h -- C-ABI of the B200 (sm_100a) compositing kernels behind MLT's plugin API.
 *
 * This is the drop-in boundary below the MLT module (mlt_b200/module/, libmltcuda.so): plain
 * pointers and sizes, no C++ types, no exceptions.  Every entry point returns 0 on success or a
 * non-zero CUDA error code (cudaError_t) / MLTCUDA_E_* value; mltcuda_last_error() gives text.
 * There is NO CPU fallback anywhere behind this header.
 *
 * Each function cites the reference (MLT 7.41.0) interface it replaces, paths relative to the
 * reference root.  All image pointers are DEVICE pointers obtained from mltcuda_malloc() unless a
 * parameter name ends in _host.  Buffers from mltcuda_malloc() are 256-byte aligned and carry 64
 * bytes of tail slack, which the vectorised kernels rely on (they issue aligned 16-byte loads
 * that may straddle the logical end of a row).
 *
 * Batching: every descriptor has `batch` (>=1; 0 is read as 1) and per-buffer `*_bstride` byte
 * strides between consecutive frames of a batch, so one launch can process many independent
 * frames (frames never interact).  A bstride of 0 re-uses the same buffer for every frame of the
 * batch (e.g. a shared luma map).
 */
#ifndef MLTCUDA_H
#define MLTCUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define MLTCUDA_ABI_VERSION 2 /* 2: mltcuda_rescale_yuv422 gained batch + strides; graphs, resize pad, mirror, opacity query, cross-context ordering added */

/* error codes beyond cudaError_t (which are < 1000) */
#define MLTCUDA_E_BADARG 10001
#define MLTCUDA_E_UNSUPPORTED 10002
#define MLTCUDA_E_NODEVICE 10003

/* Image formats: numeric values equal mlt_image_format (src/framework/mlt_types.h:40-54). */
enum {
    MLTCUDA_FMT_NONE = 0,
    MLTCUDA_FMT_RGB = 1,
    MLTCUDA_FMT_RGBA = 2,
    MLTCUDA_FMT_YUV422 = 3,
    MLTCUDA_FMT_YUV420P = 4,
    MLTCUDA_FMT_RGBA64 = 10
};

/* ------------------------------------------------------------------------------------------ */
/* Runtime: one context per (host thread, GPU).  Replaces nothing in the reference (which has no
 * device boundary); it is what mlt_consumer's real_time worker threads
 * (src/framework/mlt_consumer.c:1079-1170) bind to.                                           */

typedef struct mltcuda_ctx mltcuda_ctx;

int mltcuda_abi_version(void);
int mltcuda_device_count(void);
const char *mltcuda_last_error(void);

/* device < 0 => round-robin over visible GPUs by creation order (worker k -> GPU k mod G). */
mltcuda_ctx *mltcuda_ctx_create(int device);
void mltcuda_ctx_destroy(mltcuda_ctx *ctx);
int mltcuda_ctx_device(const mltcuda_ctx *ctx);
void *mltcuda_ctx_stream(const mltcuda_ctx *ctx); /* cudaStream_t */
int mltcuda_ctx_sync(mltcuda_ctx *ctx);
/* number of kernels launched through this context so far (bench.py's gpu_launches) */
uint64_t mltcuda_ctx_launches(const mltcuda_ctx *ctx);

/* Stream-ordered device memory (cudaMallocAsync pool; replaces mlt_pool_alloc/_release,
 * src/framework/mlt_pool.c:44,277, for device-resident images). */
void *mltcuda_malloc(mltcuda_ctx *ctx, size_t bytes);
void mltcuda_free(mltcuda_ctx *ctx, void *dev_ptr);
/* Pinned host memory for staging. */
void *mltcuda_host_alloc(size_t bytes);
void mltcuda_host_free(void *host_ptr);
/* Recycling pool of pinned buffers for images returned to host code (the replacement of
 * mlt_pool_alloc / mlt_pool_release for frames this module downloads): thread-safe, exact-size reuse. */
void *mltcuda_host_pool_alloc(size_t bytes);
void mltcuda_host_pool_release(void *host_ptr);
int mltcuda_host_register(void *host_ptr, size_t bytes);
int mltcuda_host_unregister(void *host_ptr);

/* Async copies on the context stream (cudaMemcpyAsync). */
int mltcuda_upload(mltcuda_ctx *ctx, void *dst_dev, const void *src_host, size_t bytes);
int mltcuda_download(mltcuda_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes);
/* Upload from PAGEABLE host memory -- the mlt_pool_alloc buffers stock producers hand out
 * (src/framework/mlt_pool.c:277) -- through the context's two pinned staging slots, the caller's
 * memcpy of one chunk overlapping the DMA of the previous one.  `src_host` may be released when the
 * call returns; the copy is ordered on the context stream like mltcuda_upload. */
int mltcuda_upload_staged(mltcuda_ctx *ctx, void *dst_dev, const void *src_host, size_t bytes);
int mltcuda_copy(mltcuda_ctx *ctx, void *dst_dev, const void *src_dev, size_t bytes);
int mltcuda_memset(mltcuda_ctx *ctx, void *dst_dev, int value, size_t bytes);

/* CUDA graphs.  Everything submitted to `ctx` between _begin and _end -- and to any other context that joins through
 * mltcuda_ctx_wait_event on an event recorded in between, and is joined back the same way before _end -- is
 * recorded instead of executed; mltcuda_graph_launch replays the whole sequence with one call (a launch-bound
 * per-frame loop no longer depends on the host's launch rate).  Buffers and parameters are baked in: a graph
 * replays exactly the frames it recorded.  No allocation, synchronisation or first-use initialisation may happen
 * inside the capture (run the sequence once before recording it). */
typedef struct mltcuda_graph mltcuda_graph;
int mltcuda_graph_begin(mltcuda_ctx *ctx);
mltcuda_graph *mltcuda_graph_end(mltcuda_ctx *ctx); /* NULL on error */
int mltcuda_graph_launch(mltcuda_ctx *ctx, mltcuda_graph *g);
uint64_t mltcuda_graph_kernel_nodes(const mltcuda_graph *g);
void mltcuda_graph_destroy(mltcuda_graph *g);

/* Is every byte of a device plane 0xff?  -- is_opaque() of src/modules/core/transition_luma.c:34-41 as a
 * stream-ordered query: _query enqueues the reduction and the copy of its verdict to pinned host
 * memory; after the next mltcuda_ctx_sync / event wait behind it, _result returns 1 (all 0xff) or 0.
 * Lets the host write an opaque alpha plane itself instead of copying it over PCIe. */
int mltcuda_plane_opaque_query(mltcuda_ctx *ctx, const uint8_t *plane_dev, size_t bytes);
int mltcuda_plane_opaque_result(const mltcuda_ctx *ctx);
/* The same query with the verdict in a word the CALLER owns, for contexts that more than one thread submits to
 * (_query/_result share one slot per context).  `translucent_pinned` must be pinned host memory
 * (mltcuda_host_alloc, mltcuda_host_pool_alloc or mltcuda_host_register); it is cleared by this call and, after the
 * next mltcuda_ctx_sync / event wait behind it, holds 0 if every byte of the plane is 0xff and 1 otherwise.  The
 * slot must stay allocated, and must not be reused for another query, until then. */
int mltcuda_plane_opaque_query_into(mltcuda_ctx *ctx, const uint8_t *plane_dev, size_t bytes, int *translucent_pinned);

/* CUDA-event timing on the context stream (used by bench.py; torch events do not see it). */
typedef struct mltcuda_event mltcuda_event;
mltcuda_event *mltcuda_event_create(void);
void mltcuda_event_destroy(mltcuda_event *ev);
int mltcuda_event_record(mltcuda_ctx *ctx, mltcuda_event *ev);
int mltcuda_event_sync(mltcuda_event *ev);
/* Work submitted to `ctx` after this call starts only once `ev` (recorded on any context of the
 * same process) has completed: the device-side ordering between the streams of two worker threads
 * (e.g. a frame produced on one worker's context and consumed on another's). */
int mltcuda_ctx_wait_event(mltcuda_ctx *ctx, mltcuda_event *ev);
float mltcuda_event_elapsed_ms(mltcuda_event *start, mltcuda_event *stop);
/* Work submitted to `ctx` after this call starts only once everything submitted to `other` so far
 * has completed (an event recorded on other's stream, waited on by ctx's stream; no host wait).
 * Used when an image produced on one worker's context is read or freed by another's. */
int mltcuda_ctx_wait_ctx(mltcuda_ctx *ctx, mltcuda_ctx *other);
/* Copy between buffers of two contexts, possibly on different GPUs (cudaMemcpyPeerAsync), ordered
 * after src_ctx's pending work and on dst_ctx's stream. */
int mltcuda_copy_peer(mltcuda_ctx *dst_ctx, void *dst_dev, mltcuda_ctx *src_ctx, const void *src_dev, size_t bytes);

/* ------------------------------------------------------------------------------------------ */
/* K1  imageconvert.  Replaces the cells of conversion_matrix in
 * src/modules/core/filter_imageconvert.c:423-437 (convert_yuv422_to_rgba :58-111,
 * convert_yuv422_to_rgb :113-142, convert_rgba_to_yuv422 :144-184, convert_rgb_to_yuv422
 * :186-222, convert_yuv420p_to_* :224-325, convert_yuv422_to_yuv420p :327-368,
 * convert_rgb_to_rgba :370-397, convert_rgba_to_rgb :399-419), BT.601 integer macros of
 * src/framework/mlt_frame.h:206-222.  Bit-exact.
 * rgba64 cells (NULL in the reference matrix; libswscale in avformat/filter_avcolour_space.c,
 * "parity unpinned") are defined here as: 8-bit result widened v*257; rgba64->8 bit = v>>8 for
 * colour; alpha a8<<8 up and lrint(a16/256.0) down (filter_avcolour_space.c:257-270,337-349).  */
typedef struct mltcuda_convert_desc
{
    int src_format, dst_format; /* MLTCUDA_FMT_* */
    int width, height;
    const void *src;   /* packed plane 0 (yuv420p: Y plane; U,V planes follow contiguously) */
    void *dst;         /* same convention */
    const uint8_t *src_alpha; /* optional separate alpha plane of the source frame (1 B/px) */
    uint8_t *dst_alpha;       /* alpha plane produced by rgba/rgba64 -> yuv422/rgb (required there) */
    int src_stride, dst_stride; /* bytes per row of plane 0; 0 => tightly packed */
    int batch;
    size_t src_bstride, dst_bstride, src_alpha_bstride, dst_alpha_bstride;
} mltcuda_convert_desc;

int mltcuda_convert(mltcuda_ctx *ctx, const mltcuda_convert_desc *d);

/* ------------------------------------------------------------------------------------------ */
/* K2  composite rows.  Replaces the row loop of composite_yuv
 * (src/modules/core/transition_composite.c:455-489) over composite_line_yuv (:143-182), its
 * _or/_and/_xor variants (:184-275) and the x86-64 fast path composite_line_yuv_sse2_simple
 * (src/modules/core/composite_line_yuv_sse2_simple.c:198-244).  The pointers are the ones
 * composite_yuv has computed after clipping (:407-452): `dest`/`src` already offset to the first
 * composited pixel of the first row.  Integer arithmetic, bit-exact with the x86-64 build
 * (SSE2 formula for the first width-width%8 pixels of a row when luma==NULL && width>7 and
 * op==OVER, scalar formula elsewhere).                                                         */
enum { MLTCUDA_OP_OVER = 0, MLTCUDA_OP_OR = 1, MLTCUDA_OP_AND = 2, MLTCUDA_OP_XOR = 3 };

typedef struct mltcuda_composite_desc
{
    uint8_t *dest;          /* A image (yuv422), modified in place */
    const uint8_t *src;     /* B image (yuv422) */
    const uint8_t *alpha_b; /* optional, 1 B/px, indexed like src */
    uint8_t *alpha_a;       /* optional, 1 B/px, indexed like dest, modified in place */
    const uint16_t *luma;   /* optional luma map, indexed like alpha_b */
    int width;              /* pixels per row (width_src after clipping) */
    int rows;               /* number of line_fn invocations (= ceil(height_src/step)) */
    int stride_dest, stride_src; /* BYTES between processed rows (already multiplied by field step) */
    int stride_alpha_a, stride_alpha_b; /* ELEMENTS between processed rows; luma uses stride_alpha_b */
    int weight;             /* ((1<<16)*o+50)/100, transition_composite.c:354 */
    int softness;           /* (1<<16)*softness, :353 */
    uint32_t step;          /* luma_step, :355 */
    int op;                 /* MLTCUDA_OP_* (property "operator", :1081-1095) */
    int scalar_only;        /* 1 = behave like a non-x86-64 reference build (no SSE2 groups) */
    int batch;
    size_t dest_bstride, src_bstride, alpha_a_bstride, alpha_b_bstride, luma_bstride;
} mltcuda_composite_desc;

int mltcuda_composite(mltcuda_ctx *ctx, const mltcuda_composite_desc *d);

/* Geometry front-end: the integer clipping of composite_yuv (:345-452) restated on the host so a
 * caller can hand whole frames + a rectangle.  Fills `out` (pointers offset) from whole-frame
 * pointers; returns 1 if there is nothing to composite (the reference's early returns), else 0. */
typedef struct mltcuda_composite_geometry
{
    /* struct geometry_s (:45-56) after composite_calculate/alignment */
    double x, y, w, h, o; /* item rect in normalised units, o in [0,100] */
    int nw, nh, sw, sh, x_src, y_src;
} mltcuda_composite_geometry;

int mltcuda_composite_setup(mltcuda_composite_desc *out,
                            uint8_t *p_dest, int width_dest, int height_dest,
                            const uint8_t *p_src, int width_src, int height_src,
                            const uint8_t *alpha_b, uint8_t *alpha_a,
                            const mltcuda_composite_geometry *g, int field,
                            const uint16_t *p_luma, double softness);

/* ------------------------------------------------------------------------------------------ */
/* K3  luma wipe + dissolve.  Replaces luma_composite_yuv422 / _rgba32 / _rgba64
 * (src/modules/core/transition_luma.c:327-473, 475-597, 599-723) and dissolve_slice_yuv422 +
 * composite_line_yuv_float / dissolve_slice_rgba32 / _rgba64 (:63-129, 192-292).  The opaque
 * (integer) path is bit-exact; the translucent (float) path is within +-1 LSB.  The opaque yuv422
 * dissolve is mltcuda_composite with luma=NULL (transition_luma.c:178-186).                     */
typedef struct mltcuda_luma_desc
{
    int format;              /* YUV422 / RGBA / RGBA64 */
    void *dest;              /* "a" image of luma_composite_*, modified in place */
    const void *src;         /* "b" image */
    uint8_t *alpha_dest;     /* yuv422 only, optional */
    uint8_t *alpha_src;      /* yuv422 only, optional (written when invert) */
    const uint16_t *map;     /* luma_width*luma_height uint16, NULL => dissolve */
    int luma_width, luma_height;
    int width, height;       /* *width,*height used for x_diff/y_diff (:390-391) */
    int width_dest, height_dest, width_src, height_src; /* actual image sizes */
    float pos, frame_delta, softness;
    int field_order;         /* -1 progressive, 0 lower first, 1 top first */
    int invert, alpha_over;
    int translucent;         /* -1 = decide on device like is_opaque() (:34-41), 0 opaque, 1 translucent */
    int batch;
    size_t dest_bstride, src_bstride, alpha_dest_bstride, alpha_src_bstride;
} mltcuda_luma_desc;

int mltcuda_luma(mltcuda_ctx *ctx, const mltcuda_luma_desc *d);

/* Nearest-neighbour 16-bit map scaler with optional invert: scale_luma
 * (src/modules/core/transition_composite.c:497-520). */
int mltcuda_scale_luma(mltcuda_ctx *ctx, uint16_t *dst, int dst_w, int dst_h,
                       const uint16_t *src, int src_w, int src_h, int invert_xor);

/* ------------------------------------------------------------------------------------------ */
/* K4  affine.  Replaces sliced_proc (src/modules/plus/transition_affine.c:342-409) and the
 * samplers interpNN/BL/BC_b32/_b64 (src/modules/plus/interp.h:50-334).  Mapping in double,
 * sampling in float, truncating stores; within +-1 LSB of the reference's -ffast-math build.   */
enum { MLTCUDA_INTERP_NN = 0, MLTCUDA_INTERP_BL = 1, MLTCUDA_INTERP_BC = 2 };

typedef struct mltcuda_affine_desc
{
    int format;             /* RGBA or RGBA64 (both images) */
    void *a_image;          /* modified in place */
    const void *b_image;
    int a_width, a_height, b_width, b_height;
    double matrix[3][3];    /* affine_t after get_affine + affine_scale (:616-688) */
    double lower_x, lower_y, dz, mix, x_offset, y_offset;
    double minima, xmax, ymax;
    int b_alpha;            /* is_atop */
    int interp;             /* MLTCUDA_INTERP_* */
    int batch;
    size_t a_bstride, b_bstride;
} mltcuda_affine_desc;

int mltcuda_affine(mltcuda_ctx *ctx, const mltcuda_affine_desc *d);

/* ------------------------------------------------------------------------------------------ */
/* K5  fused tractor pass.  Replaces the sequential per-transition read-modify-write passes over
 * the A frame that mlt_tractor / mlt_transition / the convert chain produce
 * (src/framework/mlt_tractor.c:348-424, src/framework/mlt_transition.c:328-381,
 * src/framework/mlt_frame.c:501-541,839-998): ONE kernel walks the op list per
 * 8-pixel group in registers and writes the output once.  Every intermediate integer step
 * (yuv<->rgb clamps, pair-averaged chroma, alpha-plane creation) is reproduced, so the result is
 * identical to running the ops one launch at a time.                                          */
enum {
    MLTCUDA_TOP_COMPOSITE = 1, /* desc.composite; dest/alpha_a ignored (the running A frame) */
    MLTCUDA_TOP_AFFINE = 2,    /* desc.affine; a_image ignored */
    MLTCUDA_TOP_LUMA = 3,      /* desc.luma (opaque integer path only); dest ignored */
    MLTCUDA_TOP_CONVERT = 4    /* to_format: YUV422 <-> RGBA of the running A frame */
};

typedef struct mltcuda_tractor_op
{
    int kind;
    int to_format; /* MLTCUDA_TOP_CONVERT: MLTCUDA_FMT_YUV422 or MLTCUDA_FMT_RGBA */
    /* MLTCUDA_TOP_COMPOSITE: fill u.composite with mltcuda_composite_setup(p_dest = desc.out,
     * width_dest = desc.width, ...): the rectangle's position inside the A frame is recovered from
     * (u.composite.dest - desc.out).  u.composite.alpha_a is ignored: whether the running A frame
     * has an alpha plane at that point follows from in_alpha and the conversions before it.
     * MLTCUDA_TOP_AFFINE: u.affine.a_image/a_width/a_height ignored (the running frame), rgba only.
     * MLTCUDA_TOP_LUMA: u.luma.dest/alpha_* ignored; opaque integer wipe, yuv422, progressive. */
    union {
        mltcuda_composite_desc composite;
        mltcuda_affine_desc affine;
        mltcuda_luma_desc luma;
    } u;
} mltcuda_tractor_op;

typedef struct mltcuda_tractor_desc
{
    int width, height;
    int in_format;            /* format of `in` : YUV422 or RGBA */
    int out_format;           /* format written to `out` */
    const void *in;           /* bottom track image (never modified) */
    const uint8_t *in_alpha;  /* optional alpha plane of the bottom track (yuv422) */
    void *out;                /* result image; may equal `in` for in-place */
    uint8_t *out_alpha;       /* optional: receives the final alpha plane if one exists */
    int n_ops;
    const mltcuda_tractor_op *ops; /* HOST array, copied into the launch */
    int batch;
    size_t in_bstride, in_alpha_bstride, out_bstride, out_alpha_bstride;
} mltcuda_tractor_desc;

#define MLTCUDA_TRACTOR_MAX_OPS 8

int mltcuda_tractor(mltcuda_ctx *ctx, const mltcuda_tractor_desc *d);

/* ------------------------------------------------------------------------------------------ */
/* Device-side synthetic producers: bit-exact with
 * src/modules/core/producer_noise.c:37-51,78-126 and producer_colour.c:60-274.               */
int mltcuda_noise(mltcuda_ctx *ctx, uint8_t *dst_yuv422, int width, int height, int position,
                  int batch, size_t dst_bstride);
/* mltcuda_noise for callers that must not allocate.  No counterpart in the reference (its producer
 * has no device memory to manage); same arguments, same bytes out as mltcuda_noise.
 *
 * Who needs it: a real_time worker thread in mid-playback, for which the device-wide lock and GPU
 * drain of a cudaMalloc/cudaFree would stall every other worker's stream; a host application that
 * has given all device memory to its own pool, or runs inside a fixed MIG/MPS memory budget, where
 * an allocation may fail at a point that cannot recover.  mltcuda_noise cudaMalloc's its skip-ahead
 * tables per context; this entry point keeps them in a static __device__ array inside the library
 * image instead (MLTCUDA_NOISE_STATIC_BYTES on every GPU that loads the library, held until the
 * process exits) and calls no cudaMalloc, cudaFree or allocating library routine, ever.
 *
 * Limits: width*height <= MLTCUDA_NOISE_STATIC_MAX_PIXELS per frame (the reach of the tables'
 * 32-bit skip-ahead) and batch <= MLTCUDA_NOISE_STATIC_MAX_BATCH; anything larger is refused with
 * MLTCUDA_E_BADARG and a message, nothing is launched.  At most 64 GPUs per process.
 *
 * First call on each GPU: looks up the array's address on that GPU and fills the tables with a
 * blocking copy on ctx's stream (thread-safe).  Make that call once per GPU at start-up, outside
 * any stream capture.  Later calls only launch the kernel.
 *
 * Sharing: all contexts of a GPU use the one array.  After the first call it is only read, so
 * calls may overlap on the GPU and nothing needs clearing between them. */
#define MLTCUDA_NOISE_STATIC_BYTES 67584 /* 2 generators x (4096 + 4096 + 256) x 4 bytes */
#define MLTCUDA_NOISE_STATIC_MAX_PIXELS 2147483647LL
#define MLTCUDA_NOISE_STATIC_MAX_BATCH 65535
int mltcuda_noise_static(mltcuda_ctx *ctx, uint8_t *dst_yuv422, int width, int height, int position,
                         int batch, size_t dst_bstride);
int mltcuda_colour(mltcuda_ctx *ctx, void *dst, uint8_t *dst_alpha, int format, int width,
                   int height, uint8_t r, uint8_t g, uint8_t b, uint8_t a);

/* Nearest-neighbour yuv422 scaler of core `rescale`: filter_scale
 * (src/modules/core/filter_rescale.c:48-120).  `batch` frames per launch (bstrides in bytes). */
int mltcuda_rescale_yuv422(mltcuda_ctx *ctx, uint8_t *dst, int dst_w, int dst_h,
                           const uint8_t *src, int src_w, int src_h,
                           int batch, size_t dst_bstride, size_t src_bstride);
/* Its alpha-plane companion scale_alpha (filter_rescale.c:122-147). */
int mltcuda_rescale_alpha(mltcuda_ctx *ctx, uint8_t *dst, int dst_w, int dst_h,
                          const uint8_t *src, int src_w, int src_h);
/* Pad / letterbox of core `resize`: resize_image (src/modules/core/filter_resize.c:65-137) for
 * format = YUV422 / RGBA / RGBA64 / RGB, and resize_alpha (:30-63) for a bare alpha plane
 * (format = MLTCUDA_FMT_NONE).  The src_w x src_h input is centred in the dst_w x dst_h output;
 * the border is black (yuv422: Y 16, UV 128) with alpha `alpha_value` (frame property
 * "resize_alpha"). */
int mltcuda_resize_pad(mltcuda_ctx *ctx, void *dst, int dst_w, int dst_h, const void *src, int src_w,
                       int src_h, int format, int alpha_value);

/* Horizontal mirror of core `mirror` in its "flip" mode (src/modules/core/filter_mirror.c:247-294): rows reversed
 * pair-wise (Y1 U Y0 V), alpha plane reversed.  Out of place; even widths (MLTCUDA_E_UNSUPPORTED otherwise). */
int mltcuda_mirror_flip(mltcuda_ctx *ctx, uint8_t *dst_yuv422, uint8_t *dst_alpha, const uint8_t *src_yuv422,
                        const uint8_t *src_alpha, int width, int height);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* MLTCUDA_H */
