/*
 * cuda_image.c -- device-resident mlt_image_private frames, per-thread contexts, the converter.
 *
 * Replaces, for frames handled by this module, the host buffers that the reference allocates with
 * mlt_pool_alloc and hands around through mlt_frame_set_image / mlt_frame_get_alpha
 * (src/framework/mlt_frame.c:335-353,501-541) and the conversion dispatch of
 * src/modules/core/filter_imageconvert.c:439-493.
 */
#include "cuda_module.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- per-thread contexts -------------------------------------------------------------------- */

#define MAX_CTX 256
static pthread_mutex_t g_ctx_lock = PTHREAD_MUTEX_INITIALIZER;
static mltcuda_ctx *g_ctx[MAX_CTX];
static int g_ctx_count = 0;
static __thread mltcuda_ctx *t_ctx = NULL;
static __thread int t_ctx_failed = 0;

mltcuda_ctx *cuda_thread_ctx(void)
{
    if (t_ctx || t_ctx_failed)
        return t_ctx;
    const char *dev = getenv("MLT_CUDA_DEVICE"); /* pin every worker to one GPU; default round-robin */
    const uint64_t st0 = CUDA_STAT_T0();
    mltcuda_ctx *c = mltcuda_ctx_create(dev && *dev ? atoi(dev) : -1);
    CUDA_STAT_END(CS_CTX_CREATE, st0);
    if (!c) {
        t_ctx_failed = 1;
        mlt_log_error(NULL, "[cuda] no usable CUDA device: %s (this module has no CPU fallback)\n", mltcuda_last_error());
        return NULL;
    }
    pthread_mutex_lock(&g_ctx_lock);
    if (g_ctx_count < MAX_CTX)
        g_ctx[g_ctx_count++] = c; /* kept until process exit: frames may outlive their worker */
    pthread_mutex_unlock(&g_ctx_lock);
    mlt_log_verbose(NULL, "[cuda] worker context %d on GPU %d\n", g_ctx_count, mltcuda_ctx_device(c));
    t_ctx = c;
    return c;
}

/* ---- MLT_CUDA_STATS ---------------------------------------------------------------------------- */

#include <stdatomic.h>
#include <time.h>
int cuda_stats_on = 0;
static _Atomic uint64_t g_stat_ns[CS_N], g_stat_calls[CS_N];
static const char *g_stat_name[CS_N] = {"composite get_image (incl. inputs)", "luma get_image (incl. inputs)", "affine get_image (incl. inputs)",
                                        "convert_image (incl. download)", "noise producer", "colour producer", "download + sync",
                                        "upload", "cuda_frame_get_image (pull input)", "worker context creation", "pinned host buffer (pool)"};
uint64_t cuda_stat_now(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t) ts.tv_sec * 1000000000ull + (uint64_t) ts.tv_nsec;
}
void cuda_stat_add(int id, uint64_t t0)
{
    atomic_fetch_add(&g_stat_ns[id], cuda_stat_now() - t0);
    atomic_fetch_add(&g_stat_calls[id], 1);
}
static void cuda_stats_print(void)
{
    for (int i = 0; i < CS_N; i++)
        if (g_stat_calls[i])
            fprintf(stderr, "[cuda stats] %-36s calls %8llu  total %9.1f ms  mean %8.1f us\n", g_stat_name[i],
                    (unsigned long long) g_stat_calls[i], g_stat_ns[i] / 1e6, g_stat_ns[i] / 1e3 / g_stat_calls[i]);
}
__attribute__((constructor)) static void cuda_stats_init(void)
{
    const char *e = getenv("MLT_CUDA_STATS");
    if (e && *e && *e != '0') {
        cuda_stats_on = 1;
        atexit(cuda_stats_print);
    }
}

int cuda_request_resident(mlt_frame frame, mlt_image_format requested)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    if (mlt_properties_get_int(p, CUDA_REQUESTED) <= 0)
        return 0;
    return requested == mlt_image_private || mlt_properties_get_int(p, CUDA_PASSTHROUGH) > 0;
}

/* ---- formats ---------------------------------------------------------------------------------- */

int cuda_format_supported(mlt_image_format f)
{
    return f == mlt_image_rgb || f == mlt_image_rgba || f == mlt_image_yuv422 || f == mlt_image_yuv420p
           || f == mlt_image_rgba64;
}

size_t cuda_format_bytes(mlt_image_format f, int w, int h)
{
    switch (f) {
    case mlt_image_rgb:
        return (size_t) w * h * 3;
    case mlt_image_rgba:
        return (size_t) w * h * 4;
    case mlt_image_yuv422:
        return (size_t) w * h * 2;
    case mlt_image_yuv420p:
        return (size_t) w * h + 2 * (size_t) (w / 2) * (h / 2);
    case mlt_image_rgba64:
        return (size_t) w * h * 8;
    default:
        return 0;
    }
}

/* ---- handles ---------------------------------------------------------------------------------- */

cuda_image cuda_image_new(mltcuda_ctx *ctx, mlt_image_format format, int width, int height, int with_alpha)
{
    size_t bytes = cuda_format_bytes(format, width, height);
    if (!ctx || !bytes)
        return NULL;
    cuda_image ci = calloc(1, sizeof(*ci));
    if (!ci)
        return NULL;
    ci->magic = CUDA_IMAGE_MAGIC, ci->ctx = ctx, ci->format = format, ci->width = width, ci->height = height;
    /* yuv420p: the reference reads up to half a chroma row past the V plane on odd lines
     * (filter_imageconvert.c:224-325); keep that read inside the allocation */
    ci->data = mltcuda_malloc(ctx, bytes + (format == mlt_image_yuv420p ? (size_t) width : 0));
    if (ci->data && with_alpha)
        ci->alpha = mltcuda_malloc(ctx, (size_t) width * height);
    if (!ci->data || (with_alpha && !ci->alpha)) {
        mlt_log_error(NULL, "[cuda] device allocation failed: %s\n", mltcuda_last_error());
        cuda_image_close(ci);
        return NULL;
    }
    return ci;
}

void cuda_image_close(cuda_image ci)
{
    if (!ci || ci->magic != CUDA_IMAGE_MAGIC)
        return;
    if (ci->data)
        mltcuda_free(ci->ctx, ci->data);
    if (ci->alpha)
        mltcuda_free(ci->ctx, ci->alpha);
    ci->magic = 0;
    free(ci);
}

int cuda_image_ensure_alpha(cuda_image ci, int fill)
{
    if (ci->alpha)
        return 0;
    ci->alpha = mltcuda_malloc(ci->ctx, (size_t) ci->width * ci->height);
    if (!ci->alpha)
        return 1;
    return mltcuda_memset(ci->ctx, ci->alpha, fill, (size_t) ci->width * ci->height);
}

void cuda_image_drop_alpha(cuda_image ci)
{
    if (ci->alpha)
        mltcuda_free(ci->ctx, ci->alpha);
    ci->alpha = NULL;
}

cuda_image cuda_frame_peek(mlt_frame frame, uint8_t *image, mlt_image_format format)
{
    if (format != mlt_image_private || !image)
        return NULL;
    const char *tag = mlt_properties_get(MLT_FRAME_PROPERTIES(frame), CUDA_IMAGE_TAG);
    if (!tag || strcmp(tag, CUDA_IMAGE_TAG_VALUE))
        return NULL;
    cuda_image ci = (cuda_image) image;
    return ci->magic == CUDA_IMAGE_MAGIC ? ci : NULL;
}

int cuda_frame_set_image(mlt_frame frame, cuda_image ci)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    mlt_properties_set_int(p, "format", mlt_image_private);
    mlt_properties_set(p, CUDA_IMAGE_TAG, CUDA_IMAGE_TAG_VALUE);
    mlt_properties_set_int(p, "width", ci->width);
    mlt_properties_set_int(p, "height", ci->height);
    mlt_frame_set_alpha(frame, NULL, 0, NULL); /* the plane lives in the handle */
    return mlt_frame_set_image(frame, (uint8_t *) ci, 0, (mlt_destructor) cuda_image_close);
}

/* ---- host <-> device ---------------------------------------------------------------------------- */

static cuda_image upload_host_image_impl(mlt_frame frame, mltcuda_ctx *ctx, const uint8_t *image, mlt_image_format format,
                                    int width, int height)
{
    const uint8_t *alpha = NULL;
    if (format != mlt_image_rgba && format != mlt_image_rgba64)
        alpha = mlt_properties_get_data(MLT_FRAME_PROPERTIES(frame), "alpha", NULL);
    cuda_image ci = cuda_image_new(ctx, format, width, height, alpha != NULL);
    if (!ci)
        return NULL;
    /* the source buffers are pageable (mlt_pool) and owned by the frame: staged through the
     * context's pinned slots, so they are free again when the call returns and the DMA runs at PCIe
     * speed, ordered on the stream ahead of the kernels that read the image */
    int rc = mltcuda_upload_staged(ctx, ci->data, image, cuda_format_bytes(format, width, height));
    if (!rc && alpha)
        rc = mltcuda_upload_staged(ctx, ci->alpha, alpha, (size_t) width * height);
    if (rc) {
        mlt_log_error(NULL, "[cuda] upload failed: %s\n", mltcuda_last_error());
        cuda_image_close(ci);
        return NULL;
    }
    return ci;
}

static cuda_image upload_host_image(mlt_frame frame, mltcuda_ctx *ctx, const uint8_t *image, mlt_image_format format, int width,
                                    int height)
{
    const uint64_t st0 = CUDA_STAT_T0();
    cuda_image ci = upload_host_image_impl(frame, ctx, image, format, width, height);
    CUDA_STAT_END(CS_UPLOAD, st0);
    return ci;
}

/* Replace the frame's private image by a host copy in ci->format (+ alpha plane). */
int cuda_download_to_frame(mlt_frame frame, cuda_image ci, uint8_t **image, mlt_image_format *format)
{
    size_t bytes = cuda_format_bytes(ci->format, ci->width, ci->height);
    size_t abytes = ci->alpha ? (size_t) ci->width * ci->height : 0;
    const uint64_t st0 = CUDA_STAT_T0();
    /* pinned destination from the recycling pool: the DMA lands in the buffer the frame will own */
    uint8_t *host = mltcuda_host_pool_alloc(bytes);
    uint8_t *ahost = abytes ? mltcuda_host_pool_alloc(abytes) : NULL;
    CUDA_STAT_END(CS_HOSTPOOL, st0);
    if (!host || (abytes && !ahost)) {
        if (host)
            mltcuda_host_pool_release(host);
        if (ahost)
            mltcuda_host_pool_release(ahost);
        return 1;
    }
    int rc = mltcuda_download(ci->ctx, host, ci->data, bytes);
    if (!rc && abytes)
        rc = mltcuda_download(ci->ctx, ahost, ci->alpha, abytes);
    if (!rc)
        rc = mltcuda_ctx_sync(ci->ctx);
    CUDA_STAT_END(CS_DOWNLOAD, st0);
    if (rc) {
        mlt_log_error(NULL, "[cuda] download failed: %s\n", mltcuda_last_error());
        if (host)
            mltcuda_host_pool_release(host);
        if (ahost)
            mltcuda_host_pool_release(ahost);
        return 1;
    }
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    mlt_image_format f = ci->format;
    mlt_properties_set(p, CUDA_IMAGE_TAG, NULL);
    mlt_properties_set_int(p, "format", f);
    mlt_frame_set_image(frame, host, (int) bytes, (mlt_destructor) mltcuda_host_pool_release); /* closes the handle */
    mlt_frame_set_alpha(frame, ahost, (int) abytes, ahost ? (mlt_destructor) mltcuda_host_pool_release : NULL);
    *image = host, *format = f;
    return 0;
}

/* ---- device conversion (imageconvert matrix) ------------------------------------------------------ */

static int matrix_cell_exists(mlt_image_format from, mlt_image_format to)
{
    /* filter_imageconvert.c:423-437 plus this module's rgba64 cells (include/mltcuda.h, K1) */
    switch (from) {
    case mlt_image_rgb:
        return to == mlt_image_rgba || to == mlt_image_yuv422;
    case mlt_image_rgba:
        return to == mlt_image_rgb || to == mlt_image_yuv422 || to == mlt_image_rgba64;
    case mlt_image_yuv422:
        return to == mlt_image_rgb || to == mlt_image_rgba || to == mlt_image_yuv420p || to == mlt_image_rgba64;
    case mlt_image_yuv420p:
        return to == mlt_image_rgb || to == mlt_image_rgba || to == mlt_image_yuv422;
    case mlt_image_rgba64:
        return to == mlt_image_rgba || to == mlt_image_yuv422;
    default:
        return 0;
    }
}

cuda_image cuda_image_convert(cuda_image src, mlt_image_format to)
{
    if (!matrix_cell_exists(src->format, to))
        return NULL;
    const int src_has_a = src->format == mlt_image_rgba || src->format == mlt_image_rgba64;
    const int dst_has_a = to == mlt_image_rgba || to == mlt_image_rgba64;
    /* alpha plane rules of convert_image (:462-480): merged into rgba output, created from rgba
     * input, otherwise carried along untouched */
    cuda_image dst = cuda_image_new(src->ctx, to, src->width, src->height, src_has_a && !dst_has_a);
    if (!dst)
        return NULL;
    mltcuda_convert_desc d;
    memset(&d, 0, sizeof(d));
    d.src_format = src->format, d.dst_format = to, d.width = src->width, d.height = src->height;
    d.src = src->data, d.dst = dst->data, d.batch = 1;
    d.src_alpha = dst_has_a ? src->alpha : NULL;
    d.dst_alpha = dst->alpha;
    if (mltcuda_convert(src->ctx, &d)) {
        mlt_log_error(NULL, "[cuda] convert %s -> %s failed: %s\n", mlt_image_format_name(src->format),
                      mlt_image_format_name(to), mltcuda_last_error());
        cuda_image_close(dst);
        return NULL;
    }
    if (!src_has_a && !dst_has_a && src->alpha) { /* carried along: move the plane */
        dst->alpha = src->alpha;
        src->alpha = NULL;
    }
    return dst;
}

/* ---- the converter ---------------------------------------------------------------------------------- */

static int cuda_convert_image_impl(mlt_frame frame, uint8_t **image, mlt_image_format *format, mlt_image_format output)
{
    if (!image || !*image)
        return 1;
    if (*format == output && *format != mlt_image_private)
        return 0;
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    /* Only an explicit request for mlt_image_private made by one of our services keeps pixels in
     * HBM.  Anything else -- in particular a CPU filter sitting between two CUDA services, which
     * asks for a host format while an outer CUDA request is still pending on the same frame --
     * gets an ordinary host image. */
    const int resident = output == mlt_image_private && mlt_properties_get_int(p, CUDA_REQUESTED) > 0;
    cuda_image ci = cuda_frame_peek(frame, *image, *format);
    if (*format == mlt_image_private && !ci)
        return 1; /* another module's private image */
    if (output == mlt_image_private && !resident)
        return 1; /* another module's request */
    if (ci && resident)
        return 0; /* stays in HBM; the asking service converts on the device if it must */
    /* the nested get_image calls below the rescale shim ask for the yuv422 the stock filter forces,
     * on behalf of a CUDA service that asked for a resident image: leave it resident */
    if (ci && output != mlt_image_private && mlt_properties_get_int(p, CUDA_PASSTHROUGH) > 0
        && mlt_properties_get_int(p, CUDA_REQUESTED) > 0)
        return 0;

    mlt_image_format target = output;
    if (!ci) { /* host image */
        if (!cuda_format_supported(*format) || (output != mlt_image_private && !cuda_format_supported(output)))
            return 1;
        if (output != mlt_image_private && !matrix_cell_exists(*format, output))
            return 1; /* same verdict as the reference: "no conversion"; let the next converter try */
        mltcuda_ctx *ctx = cuda_thread_ctx();
        if (!ctx)
            return 1;
        int w = mlt_properties_get_int(p, "width"), h = mlt_properties_get_int(p, "height");
        if (w < 1 || h < 1)
            return 1;
        ci = upload_host_image(frame, ctx, *image, *format, w, h);
        if (!ci)
            return 1;
        if (output == mlt_image_private)
            target = ci->format;
        if (ci->format != target) {
            cuda_image c2 = cuda_image_convert(ci, target);
            cuda_image_close(ci);
            if (!(ci = c2))
                return 1;
        }
        if (cuda_frame_set_image(frame, ci)) {
            cuda_image_close(ci);
            return 1;
        }
        *image = (uint8_t *) ci, *format = mlt_image_private;
        if (resident)
            return 0;
        return cuda_download_to_frame(frame, ci, image, format);
    }

    /* ours, and a host consumer is asking: convert on the device when the matrix has the cell,
     * then download; otherwise download as is and hand over to the next converter */
    if (cuda_format_supported(target) && ci->format != target && matrix_cell_exists(ci->format, target)) {
        cuda_image c2 = cuda_image_convert(ci, target);
        if (!c2)
            return 1;
        if (cuda_frame_set_image(frame, c2)) { /* closes ci */
            cuda_image_close(c2);
            return 1;
        }
        ci = c2;
    }
    if (cuda_download_to_frame(frame, ci, image, format))
        return 1;
    if (*format != output)
        return mlt_frame_next_convert_image(frame, image, format, output);
    return 0;
}

int cuda_convert_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, mlt_image_format output)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = cuda_convert_image_impl(frame, image, format, output);
    CUDA_STAT_END(CS_CONVERT, st0);
    return rc;
}

/* ---- pulling inputs ------------------------------------------------------------------------------------ */

static cuda_image cuda_frame_get_image_impl(mlt_frame frame, mlt_image_format format, int *width, int *height, int writable)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    uint8_t *img = NULL;
    mlt_image_format got = mlt_image_private; /* ask for residency explicitly (see cuda_convert_image) */
    mlt_frame_prepend_convert_image(frame, cuda_convert_image);
    mlt_properties_set_int(p, CUDA_REQUESTED, mlt_properties_get_int(p, CUDA_REQUESTED) + 1);
    mlt_properties_set_int(p, CUDA_WORK_FORMAT, format);
    int err = mlt_frame_get_image(frame, &img, &got, width, height, writable);
    mlt_properties_set_int(p, CUDA_REQUESTED, mlt_properties_get_int(p, CUDA_REQUESTED) - 1);
    if (err || !img)
        return NULL;
    cuda_image ci = cuda_frame_peek(frame, img, got);
    if (!ci) {
        /* a host image the chain could not make resident (e.g. the frame had no converter when
         * it was produced): upload it here */
        if (!cuda_format_supported(got))
            return NULL;
        mltcuda_ctx *ctx = cuda_thread_ctx();
        if (!ctx)
            return NULL;
        ci = upload_host_image(frame, ctx, img, got, *width, *height);
        if (!ci || cuda_frame_set_image(frame, ci)) {
            cuda_image_close(ci);
            return NULL;
        }
    }
    if (ci->format != format) {
        cuda_image c2 = cuda_image_convert(ci, format);
        if (!c2) {
            mlt_log_error(NULL, "[cuda] no device conversion %s -> %s\n", mlt_image_format_name(ci->format),
                          mlt_image_format_name(format));
            return NULL;
        }
        if (cuda_frame_set_image(frame, c2)) {
            cuda_image_close(c2);
            return NULL;
        }
        ci = c2;
    }
    *width = ci->width, *height = ci->height;
    return ci;
}

cuda_image cuda_frame_get_image(mlt_frame frame, mlt_image_format format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    cuda_image ci = cuda_frame_get_image_impl(frame, format, width, height, writable);
    CUDA_STAT_END(CS_PULL_A, st0);
    return ci;
}

/* ---- cached device copies of host luma maps ----------------------------------------------------------- */

typedef struct
{
    mltcuda_ctx *ctx;
    uint16_t *dev;
    const uint16_t *host;
    int w, h;
} map_cache;

static void map_cache_close(void *ptr)
{
    map_cache *m = ptr;
    if (m) {
        if (m->dev)
            mltcuda_free(m->ctx, m->dev);
        free(m);
    }
}

uint16_t *cuda_cached_map(mlt_properties owner, const char *key, mltcuda_ctx *ctx, const uint16_t *host, int w, int h)
{
    char name[96];
    snprintf(name, sizeof(name), "_cuda.%s.%p", key, (void *) ctx);
    map_cache *m = mlt_properties_get_data(owner, name, NULL);
    if (m && m->host == host && m->w == w && m->h == h)
        return m->dev;
    m = calloc(1, sizeof(*m));
    if (!m)
        return NULL;
    m->ctx = ctx, m->host = host, m->w = w, m->h = h;
    m->dev = mltcuda_malloc(ctx, (size_t) w * h * 2);
    if (!m->dev || mltcuda_upload(ctx, m->dev, host, (size_t) w * h * 2) || mltcuda_ctx_sync(ctx)) {
        map_cache_close(m);
        return NULL;
    }
    mlt_properties_set_data(owner, name, m, 0, map_cache_close, NULL); /* closes the previous one */
    return m->dev;
}
