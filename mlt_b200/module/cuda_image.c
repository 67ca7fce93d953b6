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
static mltcuda_ctx *g_ctx[MAX_CTX];      /* every context ever created (for the launch totals) */
static int g_ctx_count = 0;
static mltcuda_ctx *g_idle[MAX_CTX];     /* contexts whose worker thread has exited: re-used, never destroyed */
static int g_idle_count = 0;
static pthread_key_t g_ctx_key;
static pthread_once_t g_ctx_key_once = PTHREAD_ONCE_INIT;
static __thread mltcuda_ctx *t_ctx = NULL;
static __thread int t_ctx_failed = 0;

/* A worker thread ends (MLT consumers spawn new workers on every start / stop): its context goes back
 * to the idle list.  It is not destroyed -- frames it produced may still be queued in the consumer and
 * free their buffers through it -- but the next worker thread takes it over instead of creating another
 * stream + pinned staging, so a long-lived host (an editor restarting its consumer) does not grow. */
static void ctx_thread_exit(void *c)
{
    pthread_mutex_lock(&g_ctx_lock);
    if (c && g_idle_count < MAX_CTX)
        g_idle[g_idle_count++] = c;
    pthread_mutex_unlock(&g_ctx_lock);
}

static void ctx_key_init(void)
{
    pthread_key_create(&g_ctx_key, ctx_thread_exit);
}

mltcuda_ctx *cuda_thread_ctx(void)
{
    if (t_ctx || t_ctx_failed)
        return t_ctx;
    const char *dev = getenv("MLT_CUDA_DEVICE"); /* pin every worker to one GPU; default round-robin */
    const int want = dev && *dev ? atoi(dev) : -1;
    pthread_once(&g_ctx_key_once, ctx_key_init);
    mltcuda_ctx *c = NULL;
    pthread_mutex_lock(&g_ctx_lock);
    for (int i = g_idle_count - 1; i >= 0 && !c; i--)
        if (want < 0 || mltcuda_ctx_device(g_idle[i]) == want) {
            c = g_idle[i];
            g_idle[i] = g_idle[--g_idle_count];
        }
    pthread_mutex_unlock(&g_ctx_lock);
    if (!c) {
        const uint64_t st0 = CUDA_STAT_T0();
        c = mltcuda_ctx_create(want);
        CUDA_STAT_END(CS_CTX_CREATE, st0);
        if (!c) {
            t_ctx_failed = 1;
            mlt_log_error(NULL, "[cuda] no usable CUDA device: %s (this module has no CPU fallback)\n", mltcuda_last_error());
            return NULL;
        }
        pthread_mutex_lock(&g_ctx_lock);
        if (g_ctx_count < MAX_CTX)
            g_ctx[g_ctx_count++] = c;
        pthread_mutex_unlock(&g_ctx_lock);
        mlt_log_verbose(NULL, "[cuda] worker context %d on GPU %d\n", g_ctx_count, mltcuda_ctx_device(c));
    }
    pthread_setspecific(g_ctx_key, c);
    t_ctx = c;
    return c;
}

/* ---- MLT_CUDA_STATS ---------------------------------------------------------------------------- */

#include <stdatomic.h>
#include <time.h>
int cuda_stats_on = 0;
static _Atomic uint64_t g_stat_ns[CS_N], g_stat_calls[CS_N], g_stat_bytes[CS_N];
static const char *g_stat_name[CS_N] = {"composite get_image (incl. inputs)", "luma get_image (incl. inputs)", "affine get_image (incl. inputs)",
                                        "convert_image (incl. download)", "noise producer", "colour producer", "download + sync",
                                        "upload", "cuda_frame_get_image (pull input)", "worker context creation", "pinned host buffer (pool)",
                                        "fused pass (k_tractor launches)", "ops fused into k_tractor", "unfused composite launches",
                                        "unfused affine launches", "unfused luma launches", "unfused convert launches",
                                        "rescale launches", "resize pad launches", "mirror flip launches"};
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
void cuda_stat_count(int id, uint64_t n)
{
    atomic_fetch_add(&g_stat_calls[id], n);
}
void cuda_stat_bytes(int id, uint64_t n)
{
    atomic_fetch_add(&g_stat_bytes[id], n);
}
static void cuda_stats_print(void)
{
    for (int i = 0; i < CS_N; i++)
        if (g_stat_calls[i])
            fprintf(stderr, "[cuda stats] %-36s calls %8llu  total %9.1f ms  mean %8.1f us  bytes %llu\n", g_stat_name[i],
                    (unsigned long long) g_stat_calls[i], g_stat_ns[i] / 1e6, g_stat_ns[i] / 1e3 / g_stat_calls[i],
                    (unsigned long long) g_stat_bytes[i]);
    /* every kernel the worker contexts launched (producers included), per downloaded frame */
    unsigned long long launches = 0;
    pthread_mutex_lock(&g_ctx_lock);
    for (int i = 0; i < g_ctx_count; i++)
        launches += mltcuda_ctx_launches(g_ctx[i]);
    pthread_mutex_unlock(&g_ctx_lock);
    const unsigned long long frames = g_stat_calls[CS_DOWNLOAD];
    const unsigned long long comp = g_stat_calls[CS_FUSED] + g_stat_calls[CS_K_COMPOSITE] + g_stat_calls[CS_K_AFFINE]
                                    + g_stat_calls[CS_K_LUMA] + g_stat_calls[CS_K_CONVERT];
    if (frames)
        fprintf(stderr, "[cuda stats] kernel launches %llu over %llu downloaded frames: %.2f per frame, compositing launches per frame %.2f\n",
                launches, frames, (double) launches / frames, (double) comp / frames);
}
__attribute__((constructor)) static void cuda_stats_init(void)
{
    const char *e = getenv("MLT_CUDA_STATS");
    if (e && *e && *e != '0') {
        cuda_stats_on = 1;
        atexit(cuda_stats_print);
    }
}

static void *g_get_image[8];
static int g_get_image_n;
static pthread_mutex_t g_get_image_mu = PTHREAD_MUTEX_INITIALIZER;

void cuda_register_get_image(mlt_get_image fn)
{
    pthread_mutex_lock(&g_get_image_mu);
    int known = 0;
    for (int i = 0; i < g_get_image_n; i++)
        known |= g_get_image[i] == (void *) fn;
    if (!known && g_get_image_n < (int) (sizeof(g_get_image) / sizeof(g_get_image[0])))
        g_get_image[g_get_image_n++] = (void *) fn;
    pthread_mutex_unlock(&g_get_image_mu);
}

int cuda_is_module_get_image(void *fn)
{
    pthread_mutex_lock(&g_get_image_mu);
    int known = 0;
    for (int i = 0; fn && i < g_get_image_n; i++)
        known |= g_get_image[i] == fn;
    pthread_mutex_unlock(&g_get_image_mu);
    return known;
}

int cuda_request_resident(mlt_frame frame, mlt_image_format requested)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    if (mlt_properties_get_int(p, CUDA_REQUESTED) <= 0)
        return 0;
    return requested == mlt_image_private;
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
    ci->magic = CUDA_IMAGE_MAGIC, ci->refs = 1, ci->ctx = ctx, ci->format = format, ci->width = width, ci->height = height;
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

static void pending_release(cuda_image a, cuda_pending *p, int launched);

cuda_image cuda_image_ref(cuda_image ci)
{
    if (ci)
        __atomic_add_fetch(&ci->refs, 1, __ATOMIC_RELAXED);
    return ci;
}

void cuda_image_close(cuda_image ci)
{
    if (!ci || ci->magic != CUDA_IMAGE_MAGIC)
        return;
    if (__atomic_sub_fetch(&ci->refs, 1, __ATOMIC_ACQ_REL) > 0)
        return;
    if (ci->pending) { /* never needed: drop the list without launching */
        pending_release(ci, ci->pending, 0);
        ci->pending = NULL;
    }
    if (ci->data)
        mltcuda_free(ci->ctx, ci->data);
    if (ci->alpha)
        mltcuda_free(ci->ctx, ci->alpha);
    if (ci->ready)
        mltcuda_event_destroy(ci->ready);
    for (cuda_retired *r = ci->retired; r;) { /* every list that read them has launched and released its reference */
        cuda_retired *next = r->next;
        mltcuda_free(r->ctx, r->data);
        mltcuda_free(r->ctx, r->alpha);
        free(r);
        r = next;
    }
    ci->magic = 0;
    free(ci);
}

int cuda_image_ensure_alpha(cuda_image ci, int fill)
{
    if (cuda_image_flush(ci))
        return 1;
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

/* ---- deferred op lists: the fused tractor pass (north-star item d) ------------- */

static int fuse_enabled(void)
{
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("MLT_CUDA_FUSE"); /* 0 = launch every op on its own (the round-1 behaviour) */
        v = !(e && (!strcmp(e, "0") || !strcmp(e, "off") || !strcmp(e, "no")));
    }
    return v;
}

int cuda_image_can_defer(cuda_image a)
{
    /* what k_tractor takes (csrc/tractor.cu): whole 8-pixel groups, yuv422 or rgba running frame */
    return fuse_enabled() && a && a->width >= 8 && a->width % 8 == 0
           && (a->format == mlt_image_yuv422 || a->format == mlt_image_rgba);
}

int cuda_image_has_alpha(cuda_image a)
{
    int has = a->alpha != NULL;
    if (a->pending)
        for (int i = 0; i < a->pending->n_ops; i++)
            if (a->pending->ops[i].kind == MLTCUDA_TOP_CONVERT) /* filter_imageconvert.c:462-480 */
                has = a->pending->ops[i].to_format == MLTCUDA_FMT_YUV422;
    return has;
}

/* Drop a list: un-retain the B images.  After a launch, a B image of another context must not be
 * recycled by its own stream before the kernel that read it (on a's stream) has finished. */
static void pending_release(cuda_image a, cuda_pending *p, int launched)
{
    for (int i = 0; i < p->n_ops; i++)
        if (p->keep[i]) {
            if (launched && p->keep[i]->ctx != a->ctx)
                mltcuda_ctx_wait_ctx(p->keep[i]->ctx, a->ctx);
            for (cuda_retired *r = launched ? p->keep[i]->retired : NULL; r; r = r->next)
                if (r->ctx != a->ctx) /* the kernel may have read buffers the image has since moved out of */
                    mltcuda_ctx_wait_ctx(r->ctx, a->ctx);
            cuda_image_close(p->keep[i]);
        }
    free(p);
}

int cuda_image_localize(cuda_image b, mltcuda_ctx *ctx)
{
    if (!b || !ctx || b->ctx == ctx)
        return 0;
    if (cuda_image_flush(b))
        return 1;
    if (mltcuda_ctx_device(b->ctx) == mltcuda_ctx_device(ctx))
        return mltcuda_ctx_wait_ctx(ctx, b->ctx); /* same GPU: the pointers are valid, order the streams */
    /* another GPU (a frame cached by one worker, pulled by another): move the buffers */
    const size_t bytes = cuda_format_bytes(b->format, b->width, b->height);
    void *data = mltcuda_malloc(ctx, bytes + (b->format == mlt_image_yuv420p ? (size_t) b->width : 0));
    uint8_t *alpha = b->alpha ? mltcuda_malloc(ctx, (size_t) b->width * b->height) : NULL;
    int rc = !data || (b->alpha && !alpha);
    if (!rc)
        rc = mltcuda_copy_peer(ctx, data, b->ctx, b->data, bytes);
    if (!rc && alpha)
        rc = mltcuda_copy_peer(ctx, alpha, b->ctx, b->alpha, (size_t) b->width * b->height);
    if (rc) {
        mlt_log_error(NULL, "[cuda] moving an image from GPU %d to GPU %d failed: %s\n", mltcuda_ctx_device(b->ctx),
                      mltcuda_ctx_device(ctx), mltcuda_last_error());
        mltcuda_free(ctx, data);
        mltcuda_free(ctx, alpha);
        return 1;
    }
    mltcuda_ctx_wait_ctx(b->ctx, ctx); /* the old buffers are freed on b's stream: after the copy */
    cuda_retired *r = __atomic_load_n(&b->refs, __ATOMIC_ACQUIRE) > 1 ? calloc(1, sizeof(*r)) : NULL;
    if (r) {
        /* Somebody besides the frame holds the handle: a deferred list that has not launched yet, and that
         * recorded b->data / b->alpha when its op was appended.  Those addresses must stay good until it has run,
         * so the old buffers retire with the handle instead of being freed here. */
        r->ctx = b->ctx, r->data = b->data, r->alpha = b->alpha, r->next = b->retired;
        b->retired = r;
    } else {
        mltcuda_free(b->ctx, b->data);
        mltcuda_free(b->ctx, b->alpha);
    }
    b->data = data, b->alpha = alpha, b->ctx = ctx;
    return 0;
}

int cuda_image_mark_ready(cuda_image a)
{
    if (cuda_image_flush(a))
        return 1;
    if (!a->ready && !(a->ready = mltcuda_event_create()))
        return 1;
    return mltcuda_event_record(a->ctx, a->ready);
}

int cuda_image_wait_ready(cuda_image a)
{
    if (!a->ready)
        return cuda_image_flush(a) || mltcuda_ctx_sync(a->ctx);
    return mltcuda_event_sync(a->ready);
}

void cuda_image_used_by(cuda_image b, mltcuda_ctx *ctx)
{
    if (b && ctx && b->ctx != ctx)
        mltcuda_ctx_wait_ctx(b->ctx, ctx); /* b's buffers are recycled on b's stream: after ctx's kernel */
}

int cuda_image_defer(cuda_image a, const mltcuda_tractor_op *op, cuda_image keep)
{
    if (a->pending && a->pending->n_ops == MLTCUDA_TRACTOR_MAX_OPS && cuda_image_flush(a))
        return 1; /* a full list runs as one launch and a new list starts */
    if (keep && cuda_image_localize(keep, a->ctx))
        return 1;
    if (!a->pending) {
        a->pending = calloc(1, sizeof(*a->pending));
        if (!a->pending)
            return 1;
        a->pending->base_format = a->format;
    }
    cuda_pending *p = a->pending;
    p->ops[p->n_ops] = *op;
    p->keep[p->n_ops] = cuda_image_ref(keep);
    p->n_ops++;
    if (op->kind == MLTCUDA_TOP_CONVERT)
        a->format = (mlt_image_format) op->to_format; /* MLTCUDA_FMT_* == mlt_image_format */
    return 0;
}

static int convert_in_place(cuda_image ci, mlt_image_format to);

int cuda_image_flush(cuda_image a)
{
    cuda_pending *p = a ? a->pending : NULL;
    if (!p)
        return 0;
    a->pending = NULL;
    const mlt_image_format out_format = a->format;
    a->format = p->base_format;
    int rc = 0;
    if (p->n_ops == 1 && p->ops[0].kind == MLTCUDA_TOP_CONVERT) {
        /* a lone conversion: the dedicated K1 kernel (97 % of HBM) beats a one-op fused pass */
        rc = convert_in_place(a, out_format);
        pending_release(a, p, 1);
        return rc;
    }
    int final_alpha = a->alpha != NULL;
    for (int i = 0; i < p->n_ops; i++)
        if (p->ops[i].kind == MLTCUDA_TOP_CONVERT)
            final_alpha = p->ops[i].to_format == MLTCUDA_FMT_YUV422;
    void *out = a->data;
    uint8_t *out_alpha = a->alpha;
    if (out_format != a->format) /* same format: in place (each thread reads its group before writing it) */
        out = mltcuda_malloc(a->ctx, cuda_format_bytes(out_format, a->width, a->height));
    if (final_alpha && !out_alpha)
        out_alpha = mltcuda_malloc(a->ctx, (size_t) a->width * a->height);
    if (!out || (final_alpha && !out_alpha)) {
        mlt_log_error(NULL, "[cuda] device allocation failed: %s\n", mltcuda_last_error());
        rc = 1;
    }
    if (!rc) {
        for (int i = 0; i < p->n_ops; i++)
            if (p->ops[i].kind == MLTCUDA_TOP_COMPOSITE) /* built against CUDA_LAZY_BASE by the transition */
                p->ops[i].u.composite.dest = (uint8_t *) out + (p->ops[i].u.composite.dest - CUDA_LAZY_BASE);
        mltcuda_tractor_desc d;
        memset(&d, 0, sizeof(d));
        d.width = a->width, d.height = a->height, d.batch = 1;
        d.in_format = a->format, d.out_format = out_format;
        d.in = a->data, d.in_alpha = a->format == mlt_image_yuv422 ? a->alpha : NULL;
        d.out = out, d.out_alpha = final_alpha ? out_alpha : NULL;
        d.n_ops = p->n_ops, d.ops = p->ops;
        rc = mltcuda_tractor(a->ctx, &d);
        if (rc)
            mlt_log_error(NULL, "[cuda] fused tractor pass (%d ops) failed: %s\n", p->n_ops, mltcuda_last_error());
        CUDA_STAT_COUNT(CS_FUSED, 1);
        CUDA_STAT_COUNT(CS_FUSED_OPS, p->n_ops);
    }
    if (out && out != a->data) {
        mltcuda_free(a->ctx, a->data); /* stream-ordered: after the launch that read it */
        a->data = out;
    }
    if (final_alpha) {
        if (out_alpha)
            a->alpha = out_alpha;
    } else if (a->alpha) { /* merged into the rgba pixels (filter_imageconvert.c:475-477) */
        mltcuda_free(a->ctx, a->alpha);
        a->alpha = NULL;
    }
    if (out)
        a->format = out_format;
    pending_release(a, p, 1);
    return rc;
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
    if (cuda_stats_on && ci)
        cuda_stat_bytes(CS_UPLOAD, cuda_format_bytes(format, width, height) + (ci->alpha ? (size_t) width * height : 0));
    return ci;
}

/* Replace the frame's private image by a host copy in ci->format (+ alpha plane). */
int cuda_download_to_frame(mlt_frame frame, cuda_image ci, uint8_t **image, mlt_image_format *format)
{
    if (cuda_image_flush(ci)) /* the deferred op list runs here: one fused launch, then the DMA */
        return 1;
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
    /* An alpha plane that is 0xff everywhere (the usual case: the rgba -> yuv422 conversion of an opaque
     * picture creates one, filter_imageconvert.c:144-184) does not cross PCIe: the device answers
     * is_opaque() behind the image copy and the host writes the plane itself (a third of the bytes of a
     * yuv422 frame).  MLT_CUDA_ALPHA_DMA=1 forces the copy. */
    static int alpha_dma = -1;
    if (alpha_dma < 0) {
        const char *e = getenv("MLT_CUDA_ALPHA_DMA");
        alpha_dma = e && *e && *e != '0';
    }
    int rc = mltcuda_download(ci->ctx, host, ci->data, bytes);
    /* The verdict lands in a pinned word of this call's own: ci->ctx may be another thread's context (an adopted
     * idle one, an image cloned onto its producer's context, a consumer asking again), and two downloads on one
     * context must not read each other's answer. */
    int *translucent = !rc && abytes && !alpha_dma && !((uintptr_t) ci->alpha & 15) ? mltcuda_host_pool_alloc(sizeof(int)) : NULL;
    const int query = translucent != NULL;
    if (query)
        rc = mltcuda_plane_opaque_query_into(ci->ctx, ci->alpha, abytes, translucent);
    else if (!rc && abytes)
        rc = mltcuda_download(ci->ctx, ahost, ci->alpha, abytes);
    if (!rc)
        rc = mltcuda_ctx_sync(ci->ctx);
    size_t moved = bytes + abytes;
    if (!rc && query) {
        if (!*translucent) {
            memset(ahost, 0xff, abytes);
            moved = bytes;
        } else {
            rc = mltcuda_download(ci->ctx, ahost, ci->alpha, abytes);
            if (!rc)
                rc = mltcuda_ctx_sync(ci->ctx);
        }
    }
    if (translucent)
        mltcuda_host_pool_release(translucent);
    CUDA_STAT_END(CS_DOWNLOAD, st0);
    if (cuda_stats_on)
        cuda_stat_bytes(CS_DOWNLOAD, moved);
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

/* the two cells the fused pass carries in registers */
static int fusable_pair(mlt_image_format from, mlt_image_format to)
{
    return (from == mlt_image_yuv422 && to == mlt_image_rgba) || (from == mlt_image_rgba && to == mlt_image_yuv422);
}

cuda_image cuda_image_convert(cuda_image src, mlt_image_format to)
{
    if (!matrix_cell_exists(src->format, to) || cuda_image_flush(src))
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
    CUDA_STAT_COUNT(CS_K_CONVERT, 1);
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

/* Same, keeping the handle: the buffers of `ci` are replaced by the converted ones. */
static int convert_in_place(cuda_image ci, mlt_image_format to)
{
    if (ci->format == to)
        return 0;
    cuda_image c2 = cuda_image_convert(ci, to);
    if (!c2)
        return 1;
    mltcuda_free(ci->ctx, ci->data);
    mltcuda_free(ci->ctx, ci->alpha);
    ci->data = c2->data, ci->alpha = c2->alpha, ci->format = to;
    c2->data = NULL, c2->alpha = NULL;
    cuda_image_close(c2);
    return 0;
}

/* Bring the frame's image `*pci` to `format`.  lazy: the conversion may join / start the deferred
 * list (A role).  A handle that some pending list still reads (refs > 1) is never rewritten: the
 * frame gets a new image and the old one lives on for its readers. */
static int frame_image_to_format(mlt_frame frame, cuda_image *pci, mlt_image_format format, int lazy)
{
    cuda_image ci = *pci;
    if (ci->format == format)
        return lazy ? 0 : cuda_image_flush(ci);
    const int shared = __atomic_load_n(&ci->refs, __ATOMIC_ACQUIRE) > 1;
    if (!shared && fusable_pair(ci->format, format) && cuda_image_can_defer(ci) && (lazy || ci->pending)) {
        mltcuda_tractor_op op;
        memset(&op, 0, sizeof(op));
        op.kind = MLTCUDA_TOP_CONVERT, op.to_format = format;
        if (cuda_image_defer(ci, &op, NULL))
            return 1;
        return lazy ? 0 : cuda_image_flush(ci);
    }
    if (shared) {
        cuda_image c2 = cuda_image_convert(ci, format);
        if (c2 && ci->alpha == NULL && c2->alpha && ci->format != mlt_image_rgba && ci->format != mlt_image_rgba64) {
            /* cuda_image_convert MOVED a carried-along alpha plane out of the shared source (the only way a source
             * without alpha in its pixels yields a plane, whatever the target -- rgb / yuv420p -> yuv422 included):
             * give it a copy back */
            ci->alpha = mltcuda_malloc(ci->ctx, (size_t) ci->width * ci->height);
            if (ci->alpha)
                mltcuda_copy(ci->ctx, ci->alpha, c2->alpha, (size_t) ci->width * ci->height);
        }
        if (!c2 || cuda_frame_set_image(frame, c2)) {
            cuda_image_close(c2);
            return 1;
        }
        *pci = c2;
        return 0;
    }
    if (convert_in_place(ci, format)) {
        mlt_log_error(NULL, "[cuda] no device conversion %s -> %s\n", mlt_image_format_name(ci->format),
                      mlt_image_format_name(format));
        return 1;
    }
    return 0;
}

/* ---- the converters ---------------------------------------------------------------------------------- */

/* Is the converter that would run after the current one this module's imageconvert (or nothing)?
 * Then converting on the device before the download gives what the chain would have produced;
 * a foreign converter ahead of it (avcolor_space in a full MLT build, core/loader.ini:10) keeps its
 * first refusal, as in the reference. */
static int cuda_imageconvert_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, mlt_image_format output);
static int next_converter_is_ours(mlt_frame frame)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    mlt_deque list = mlt_properties_get_data(p, "_convert_image_callbacks", NULL); /* mlt_frame.c:835 */
    if (!list)
        return 1;
    for (int i = mlt_properties_get_int(p, "_convert_image_idx") + 1; i < mlt_deque_count(list); i++) {
        void *fn = mlt_deque_peek(list, i);
        if (fn == (void *) cuda_imageconvert_image || fn == (void *) cuda_convert_image)
            return 1;
        return 0;
    }
    return 1;
}

/* full = the `imageconvert` filter's callback (every cell of the matrix, host images included);
 * otherwise the residency converter the CUDA services prepend (device images only). */
static int convert_image_impl(mlt_frame frame, uint8_t **image, mlt_image_format *format, mlt_image_format output, int full)
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
    /* Device-consumer hand-off (consumer_null.c): the consumer marked THIS frame for its own, next request, so
     * the host format it nominally asks for (mlt_image_private cannot be named in "mlt_image_format": it sorts
     * after mlt_image_invalid, mlt_types.h:52-53) is answered with the resident picture -- the deferred list
     * runs here, in the worker thread, an event is recorded for the consumer, nothing is copied.  The mark is
     * spent here: whoever asks this frame for a host format afterwards gets pixels. */
    if (ci && mlt_properties_get_int(p, CUDA_HANDOFF) > 0 && mlt_properties_get_int(p, CUDA_REQUESTED) <= 0) {
        mlt_properties_set_int(p, CUDA_HANDOFF, 0);
        return cuda_image_mark_ready(ci);
    }

    if (!ci) { /* host image */
        /* host -> host is the imageconvert filter's job, in the filter's place in the chain (the
         * reference appends it, filter_imageconvert.c:498-502: avcolor_space gets the first try) */
        if (output != mlt_image_private && !full)
            return 1;
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
        if (output != mlt_image_private && convert_in_place(ci, output)) {
            cuda_image_close(ci);
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

    /* ours, and a host consumer is asking: convert on the device when the matrix has the cell and
     * the chain would have ended in imageconvert anyway, then download; otherwise download as is
     * and hand over to the next converter */
    if (cuda_format_supported(output) && ci->format != output && matrix_cell_exists(ci->format, output)
        && (full || next_converter_is_ours(frame))) {
        if (frame_image_to_format(frame, &ci, output, 1)) /* joins the deferred list when it can: the download flushes */
            return 1;
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
    int rc = convert_image_impl(frame, image, format, output, 0);
    CUDA_STAT_END(CS_CONVERT, st0);
    return rc;
}

static int cuda_imageconvert_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, mlt_image_format output)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = convert_image_impl(frame, image, format, output, 1);
    CUDA_STAT_END(CS_CONVERT, st0);
    return rc;
}

/* filter_imageconvert.c (this module): appended, like the reference's filter_process (:498-502) */
void cuda_imageconvert_attach(mlt_frame frame)
{
    mlt_frame_append_convert_image(frame, cuda_imageconvert_image);
}

/* ---- pulling inputs ------------------------------------------------------------------------------------ */

static cuda_image cuda_frame_get_image_impl(mlt_frame frame, mlt_image_format format, int *width, int *height, int writable,
                                            int lazy)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    uint8_t *img = NULL;
    mlt_image_format got = mlt_image_private; /* ask for residency explicitly (see cuda_convert_image) */
    mlt_frame_prepend_convert_image(frame, cuda_convert_image);
    const int saved_work = mlt_properties_get_int(p, CUDA_WORK_FORMAT);
    mlt_properties_set_int(p, CUDA_REQUESTED, mlt_properties_get_int(p, CUDA_REQUESTED) + 1);
    mlt_properties_set_int(p, CUDA_WORK_FORMAT, format);
    int err = mlt_frame_get_image(frame, &img, &got, width, height, writable);
    mlt_properties_set_int(p, CUDA_REQUESTED, mlt_properties_get_int(p, CUDA_REQUESTED) - 1);
    mlt_properties_set_int(p, CUDA_WORK_FORMAT, saved_work);
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
    if (lazy && __atomic_load_n(&ci->refs, __ATOMIC_ACQUIRE) > 1) {
        /* a deferred list somewhere still reads this handle as its B image: the A role writes, so
         * it gets its own copy */
        if (cuda_image_flush(ci))
            return NULL;
        cuda_image c2 = cuda_image_new(ci->ctx, ci->format, ci->width, ci->height, ci->alpha != NULL);
        if (!c2 || mltcuda_copy(ci->ctx, c2->data, ci->data, cuda_format_bytes(ci->format, ci->width, ci->height))
            || (ci->alpha && mltcuda_copy(ci->ctx, c2->alpha, ci->alpha, (size_t) ci->width * ci->height))
            || cuda_frame_set_image(frame, c2)) {
            cuda_image_close(c2);
            return NULL;
        }
        ci = c2;
    }
    if (frame_image_to_format(frame, &ci, format, lazy))
        return NULL;
    *width = ci->width, *height = ci->height;
    return ci;
}

cuda_image cuda_frame_get_image(mlt_frame frame, mlt_image_format format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    cuda_image ci = cuda_frame_get_image_impl(frame, format, width, height, writable, 0);
    CUDA_STAT_END(CS_PULL_A, st0);
    return ci;
}

cuda_image cuda_frame_get_image_a(mlt_frame frame, mlt_image_format format, int *width, int *height)
{
    const uint64_t st0 = CUDA_STAT_T0();
    cuda_image ci = cuda_frame_get_image_impl(frame, format, width, height, 1, 1);
    CUDA_STAT_END(CS_PULL_A, st0);
    return ci;
}

/* ---- cached device copies of host luma maps ----------------------------------------------------------- */

typedef struct
{
    mltcuda_ctx *ctx;
    uint16_t *dev;
    const uint16_t *host;
    int w, h, generation;
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

/* `generation` counts the replacements of the host map (the owner bumps it whenever it installs a new
 * bitmap): mlt_pool recycles same-size blocks, so pointer + size alone cannot tell a regenerated map
 * from the one a context uploaded earlier. */
uint16_t *cuda_cached_map(mlt_properties owner, const char *key, mltcuda_ctx *ctx, const uint16_t *host, int w, int h, int generation)
{
    char name[96];
    snprintf(name, sizeof(name), "_cuda.%s.%p", key, (void *) ctx);
    map_cache *m = mlt_properties_get_data(owner, name, NULL);
    if (m && m->host == host && m->w == w && m->h == h && m->generation == generation)
        return m->dev;
    m = calloc(1, sizeof(*m));
    if (!m)
        return NULL;
    m->ctx = ctx, m->host = host, m->w = w, m->h = h, m->generation = generation;
    m->dev = mltcuda_malloc(ctx, (size_t) w * h * 2);
    if (!m->dev || mltcuda_upload(ctx, m->dev, host, (size_t) w * h * 2) || mltcuda_ctx_sync(ctx)) {
        map_cache_close(m);
        return NULL;
    }
    mlt_properties_set_data(owner, name, m, 0, map_cache_close, NULL); /* closes the previous one */
    return m->dev;
}
