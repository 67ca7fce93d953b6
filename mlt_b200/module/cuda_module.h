/*
 * cuda_module.h -- shared declarations of the MLT "cuda" module (libmltcuda.so).
 *
 * The module is the host side (plain C, links libmlt-7) of the B200 compositing path: it registers
 * the services `composite`, `luma`, `affine` (transitions) and `imageconvert` (filter) under the
 * reference's names and properties, and runs their pixel work through the C-ABI of
 * include/mltcuda.h (libmltcuda_rt.so).  There is no CPU pixel path in this module.
 *
 * Device residency protocol (precedent: src/modules/placebo/gpu_context.c:403-447 in the
 * reference): a frame whose pixels live in HBM carries format mlt_image_private, the image pointer
 * is a `cuda_image` handle owned by the frame (destructor frees the device buffers) and the frame
 * property "mlt_image_private" is the string "cuda".  A CUDA service that pulls an input raises the
 * frame's "_cuda_requested" counter around mlt_frame_get_image(); while it is raised the module's
 * converter (prepended to the frame's convert chain) leaves private images alone and turns host
 * images into private ones, so chained CUDA services never round-trip through host memory.  When
 * the counter is zero (a CPU filter or the consumer is asking) the converter downloads.
 */
#ifndef MLT_CUDA_MODULE_H
#define MLT_CUDA_MODULE_H

#include <framework/mlt.h>
#include <framework/mlt_luma_map.h>
#include <stdint.h>

#include "mltcuda.h"

#define CUDA_IMAGE_MAGIC 0x43554441u /* "CUDA" */
#define CUDA_IMAGE_TAG "mlt_image_private"
#define CUDA_IMAGE_TAG_VALUE "cuda"
#define CUDA_REQUESTED "_cuda_requested"
#define CUDA_CONSUMER "_cuda_consumer" /* set by a device consumer (consumer_null.c) on the frames it handed off: their image is a cuda_image */
/* One-shot, set by a device consumer on the ONE frame whose next host-format request provably comes from the
 * consumer itself (consumer_null.c: consumer_is_sole_requester): the converter answers that request with the
 * resident picture instead of a download, and clears the mark.  Deliberately without the "consumer." prefix:
 * mlt_tractor (mlt_tractor.c:366) and mlt_transition (mlt_transition.c:378) copy "consumer." properties to track
 * and B frames, where a stock CPU service asking for yuv422 must get pixels, not a handle. */
#define CUDA_HANDOFF "_cuda_handoff"
#define CUDA_WORK_FORMAT "_cuda_work_format" /* format the pulling service will work in (hint for device producers) */

/* Deferred op list of one image (precedent: the per-frame effect chain of
 * src/modules/movit/filter_movit_convert.cpp:623-690).  While a CUDA service is the puller, the
 * transitions and the converter do not launch: they append their op to the A image's list, and the
 * list runs as ONE fused k_tractor launch (mltcuda_tractor) when somebody needs the pixels -- the
 * converter that sees a host request, a service that uses the image as its B input, or an op the
 * fused kernel refuses (those flush, then launch unfused). */
typedef struct cuda_image_s *cuda_image;
typedef struct cuda_pending_s
{
    mlt_image_format base_format; /* format of the buffers in `data` / `alpha` (the bottom track) */
    int n_ops;
    mltcuda_tractor_op ops[MLTCUDA_TRACTOR_MAX_OPS];
    cuda_image keep[MLTCUDA_TRACTOR_MAX_OPS]; /* B images the ops read: retained until the launch */
} cuda_pending;

/* Buffers an image had before it moved to another GPU while a deferred list still held their addresses
 * (cuda_image_localize): kept until the last reference to the handle goes. */
typedef struct cuda_retired_s
{
    mltcuda_ctx *ctx;
    void *data, *alpha;
    struct cuda_retired_s *next;
} cuda_retired;

struct cuda_image_s
{
    uint32_t magic;
    int refs;              /* owners: the frame that carries the handle + pending lists that read it */
    mltcuda_ctx *ctx;      /* context (GPU + stream + pool) the buffers belong to */
    mlt_image_format format; /* pixel format once the pending ops have run: rgb, rgba, yuv422, yuv420p, rgba64 */
    int width, height;
    void *data;            /* device; with pending ops: the bottom track in pending->base_format */
    uint8_t *alpha;        /* device, 1 B/px, or NULL (same remark) */
    cuda_pending *pending; /* NULL = the buffers hold the image */
    mltcuda_event *ready;  /* recorded behind the last launch when the image is handed to a device consumer */
    cuda_retired *retired; /* earlier buffers that pending lists of other images may still read */
};

/* --- runtime (cuda_image.c) -------------------------------------------------------------- */

/* The calling thread's context; created on first use, worker k -> GPU k mod G.
 * NULL (and an error logged) when no device is usable: callers then fail their get_image. */
mltcuda_ctx *cuda_thread_ctx(void);

/* MLT_CUDA_STATS=1: wall time per stage, summed over the worker threads, printed at exit (stderr).
 * Inclusive times: a transition's figure contains the get_image of its inputs. */
enum { CS_COMPOSITE, CS_LUMA, CS_AFFINE, CS_CONVERT, CS_NOISE, CS_COLOUR, CS_DOWNLOAD, CS_UPLOAD, CS_PULL_A, CS_CTX_CREATE, CS_HOSTPOOL,
       CS_FUSED, CS_FUSED_OPS, CS_K_COMPOSITE, CS_K_AFFINE, CS_K_LUMA, CS_K_CONVERT, CS_K_RESCALE, CS_K_PAD, CS_K_MIRROR, CS_N };
/* count-only entries (no time): */
void cuda_stat_count(int id, uint64_t n);
void cuda_stat_bytes(int id, uint64_t n);
#define CUDA_STAT_COUNT(id, n) \
    do { \
        if (cuda_stats_on) \
            cuda_stat_count(id, n); \
    } while (0)
extern int cuda_stats_on;
uint64_t cuda_stat_now(void);
void cuda_stat_add(int id, uint64_t t0);
#define CUDA_STAT_T0() (cuda_stats_on ? cuda_stat_now() : 0)
#define CUDA_STAT_END(id, t0) \
    do { \
        if (cuda_stats_on) \
            cuda_stat_add(id, t0); \
    } while (0)


size_t cuda_format_bytes(mlt_image_format format, int width, int height);
int cuda_format_supported(mlt_image_format format);

cuda_image cuda_image_new(mltcuda_ctx *ctx, mlt_image_format format, int width, int height, int with_alpha);
void cuda_image_close(cuda_image ci); /* drops one reference */
cuda_image cuda_image_ref(cuda_image ci);

/* --- deferred op lists (cuda_image.c) -------------------------------------------------------- */
/* Composite ops of a deferred list are built by mltcuda_composite_setup() against this fake frame
 * pointer; cuda_image_flush() rebases them onto the output buffer. */
#define CUDA_LAZY_BASE ((uint8_t *) (uintptr_t) 0x400000000000ull)
/* May ops on `a` be deferred?  (MLT_CUDA_FUSE != 0, yuv422 / rgba, width % 8 == 0: what k_tractor takes) */
int cuda_image_can_defer(cuda_image a);
/* Append `op` (reading B image `keep`, may be NULL) to a's list.  0 on success. */
int cuda_image_defer(cuda_image a, const mltcuda_tractor_op *op, cuda_image keep);
/* Run a's pending ops (one fused launch) so that data / alpha hold the image.  0 on success. */
int cuda_image_flush(cuda_image a);
/* Device-consumer hand-off: run the pending ops and record `ready` behind them / wait for it. */
int cuda_image_mark_ready(cuda_image a);
int cuda_image_wait_ready(cuda_image a);
/* Will the image have an alpha plane once its pending ops have run? */
int cuda_image_has_alpha(cuda_image a);
/* Make `b` (materialised) readable by kernels launched on `ctx`: same context -> nothing; same GPU,
 * other stream -> stream ordering; other GPU -> the buffers move to `ctx`.  0 on success. */
int cuda_image_localize(cuda_image b, mltcuda_ctx *ctx);
/* ... and after a kernel on `ctx` has read it: b's own stream must not recycle the buffers earlier. */
void cuda_image_used_by(cuda_image b, mltcuda_ctx *ctx);
int cuda_image_ensure_alpha(cuda_image ci, int fill); /* allocates + memsets if missing */
void cuda_image_drop_alpha(cuda_image ci);

/* Is the frame's current image one of ours? */
cuda_image cuda_frame_peek(mlt_frame frame, uint8_t *image, mlt_image_format format);
/* Install `ci` as the frame's image (frame takes ownership). */
int cuda_frame_set_image(mlt_frame frame, cuda_image ci);

/* Pull the frame's image as a device image in `format` (yuv422 / rgba / rgba64) at the requested
 * size (in/out like mlt_frame_get_image).  Uploads / converts on the device as needed.  The
 * returned handle stays owned by the frame and is materialised (no pending ops): the B role.
 * NULL on error. */
cuda_image cuda_frame_get_image(mlt_frame frame, mlt_image_format format, int *width, int *height, int writable);
/* The A role: the image a transition composites INTO.  May come back with pending ops (the
 * conversion to `format` is itself deferred when it can be) and is exclusively owned by the frame
 * (a handle that a pending list still reads is cloned first). */
cuda_image cuda_frame_get_image_a(mlt_frame frame, mlt_image_format format, int *width, int *height);

/* Replace the frame's device image `ci` by a host copy (pinned, recycled) in ci->format, with its
 * alpha plane; *image / *format are updated.  Closes the handle. */
int cuda_download_to_frame(mlt_frame frame, cuda_image ci, uint8_t **image, mlt_image_format *format);

/* Convert a device image to another format on the device (imageconvert matrix).  Returns a NEW
 * image (caller owns) or NULL if the cell does not exist / on error. */
cuda_image cuda_image_convert(cuda_image src, mlt_image_format to);

/* The converter callback (mlt_convert_image) the services prepend to frames. */
int cuda_convert_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, mlt_image_format output);

/* filter_rescale.c: the loader's stock `rescale` / `resize` normalisers answer CUDA requests on the device */
void cuda_rescale_install(mlt_filter stock_rescale);
void cuda_resize_install(mlt_filter stock_resize);
void cuda_mirror_install(mlt_filter stock_mirror); /* mirror:flip between a device producer and a CUDA transition */
/* cuda_image.c: registers the imageconvert filter's converter on a frame (appended, like the reference) */
void cuda_imageconvert_attach(mlt_frame frame);
/* The get_image callbacks this module's transitions push (registered by their constructors): lets a device consumer
 * check that the service on top of a frame's image stack is one that understands the hand-off. */
void cuda_register_get_image(mlt_get_image fn);
int cuda_is_module_get_image(void *fn);
/* true if a device producer may answer this request with a resident image */
int cuda_request_resident(mlt_frame frame, mlt_image_format requested);

/* Device copy of a host uint16 luma map, cached per context on `owner` under `key` until the host pointer,
 * the size or `generation` (bumped by the owner whenever it replaces the map) changes. */
#define CUDA_MAP_GENERATION "_cuda_map_generation"
uint16_t *cuda_cached_map(mlt_properties owner, const char *key, mltcuda_ctx *ctx, const uint16_t *host, int w, int h, int generation);

/* --- constructors (one file per service) ---------------------------------------------------- */
mlt_producer cuda_producer_noise_init(mlt_profile, mlt_service_type, const char *, char *);
mlt_producer cuda_producer_colour_init(mlt_profile, mlt_service_type, const char *, char *);
extern __thread int cuda_creating_stock_producer;
mlt_transition cuda_transition_composite_init(mlt_profile, mlt_service_type, const char *, char *);
mlt_transition cuda_transition_luma_init(mlt_profile, mlt_service_type, const char *, char *);
mlt_transition cuda_transition_affine_init(mlt_profile, mlt_service_type, const char *, char *);
mlt_filter cuda_filter_imageconvert_init(mlt_profile, mlt_service_type, const char *, char *);
mlt_consumer cuda_consumer_null_init(mlt_profile, mlt_service_type, const char *, char *);

#endif
