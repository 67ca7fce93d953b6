/*
 * transition_luma.c (cuda module) -- the `luma` wipe / dissolve transition on the GPU.
 *
 * Drop-in for src/modules/core/transition_luma.c of the reference: same service name and
 * properties (resource, softness, reverse, invert, producer.*, alpha_over, fix_background_alpha,
 * fixed, progressive, factory).  The map is loaded on the host exactly as the reference's
 * transition_get_image does (:770-936: PGM, generated pattern, or a producer frame), uploaded once
 * per GPU and cached; the pixel work is mltcuda_luma (K3) on device-resident images.
 */
#include "cuda_module.h"

#include <limits.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* yuv422_to_luma16 (:725-744): luma of a clip frame -> 16-bit map */
static uint16_t *clip_luma16(const uint8_t *image, int width, int height, int full_range)
{
    uint16_t *map = mlt_pool_alloc((size_t) width * height * sizeof(uint16_t));
    if (!map)
        return NULL;
    int lo = full_range ? 0 : 16, hi = full_range ? 255 : 235 - lo, k = full_range ? 256 : 299;
    for (int i = 0; i < width * height; i++)
        map[i] = CLAMP(image[i << 1] - lo, 0, hi) * k;
    return map;
}

/* Refresh the cached host map when "resource" changed.  Caller holds the service lock.
 * Returns the producer when the map follows a video clip (map regenerated every frame). */
static mlt_producer refresh_map(mlt_transition self, mlt_frame a_frame, uint16_t **bitmap, int *lw, int *lh)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_profile profile = mlt_service_profile(MLT_TRANSITION_SERVICE(self));
    char *current = mlt_properties_get(props, "_resource");
    mlt_producer producer = mlt_properties_get_data(props, "producer", NULL);
    char *resource = mlt_properties_get(props, "resource"), *named = resource;
    char path[PATH_MAX];

    if (!resource || !(producer || !current || strcmp(resource, current)))
        return producer;
    char *ext = strrchr(resource, '.');
    if (strchr(resource, '%')) {
        snprintf(path, sizeof(path) - 5, "%s/lumas/%s/%s", mlt_environment("MLT_DATA"), mlt_profile_lumas_dir(profile),
                 strchr(resource, '%') + 1);
        FILE *t = mlt_fopen(path, "r");
        if (!t) {
            strcat(path, ".png");
            t = mlt_fopen(path, "r");
        }
        if (t) {
            fclose(t);
            resource = path;
        }
        ext = strrchr(resource, '.');
    }
    if (ext && !strcmp(ext, ".pgm")) {
        *bitmap = NULL;
        if (mlt_luma_map_from_pgm(resource, bitmap, lw, lh)) {
            mlt_luma_map gen = mlt_luma_map_new(named);
            if (profile)
                gen->w = profile->width, gen->h = profile->height;
            *bitmap = mlt_luma_map_render(gen);
            *lw = gen->w, *lh = gen->h;
            free(gen);
        }
        mlt_properties_set_int(props, "width", *lw);
        mlt_properties_set_int(props, "height", *lh);
        mlt_properties_set(props, "_resource", named);
        mlt_properties_set_int(props, CUDA_MAP_GENERATION, mlt_properties_get_int(props, CUDA_MAP_GENERATION) + 1);
        mlt_properties_set_data(props, "bitmap", *bitmap, *lw * *lh * 2, mlt_pool_release, NULL);
        mlt_properties_clear(props, "producer");
        return NULL;
    }
    if (!*resource) {
        *bitmap = NULL;
        mlt_properties_set(props, "_resource", NULL);
        mlt_properties_set_int(props, CUDA_MAP_GENERATION, mlt_properties_get_int(props, CUDA_MAP_GENERATION) + 1);
        mlt_properties_set_data(props, "bitmap", NULL, 0, mlt_pool_release, NULL);
        mlt_properties_clear(props, "producer");
        return NULL;
    }
    if (!producer || !current || strcmp(resource, current)) {
        producer = mlt_factory_producer(profile, mlt_properties_get(props, "factory"), resource);
        if (producer)
            mlt_properties_set(props, "_resource", resource);
    }
    if (!producer)
        return NULL;
    mlt_properties pp = MLT_PRODUCER_PROPERTIES(producer);
    mlt_properties_set(pp, "eof", "loop");
    mlt_properties_pass(pp, props, "producer.");
    const char *svc = mlt_properties_get(pp, "mlt_service");
    int is_clip = mlt_producer_get_length(producer) > 1 && mlt_properties_get_int(pp, "video_index") >= 0 && svc
                  && !strncmp("avformat", svc, 8);
    if (is_clip) {
        if (!mlt_properties_get(pp, "producer.eof"))
            mlt_properties_set(pp, "eof", "pause");
        mlt_producer_seek(producer, mlt_transition_get_position(self, a_frame));
    }
    mlt_frame lf = NULL;
    if (mlt_service_get_frame(MLT_PRODUCER_SERVICE(producer), &lf, 0) == 0) {
        uint8_t *img = NULL;
        mlt_image_format f = mlt_image_yuv422;
        mlt_frame_get_image(lf, &img, &f, lw, lh, 0);
        if (img) {
            if (is_clip)
                *bitmap = clip_luma16(img, *lw, *lh, mlt_properties_get_int(MLT_FRAME_PROPERTIES(lf), "full_range"));
            else
                mlt_luma_map_from_yuv422(img, bitmap, *lw, *lh);
        }
        mlt_properties_set_int(props, "width", *lw);
        mlt_properties_set_int(props, "height", *lh);
        mlt_properties_set_int(props, CUDA_MAP_GENERATION, mlt_properties_get_int(props, CUDA_MAP_GENERATION) + 1);
        mlt_properties_set_data(props, "bitmap", *bitmap, *lw * *lh * 2, mlt_pool_release, NULL);
        mlt_frame_close(lf);
    }
    if (is_clip) {
        mlt_properties_set_data(props, "producer", producer, 0, (mlt_destructor) mlt_producer_close, NULL);
        return producer;
    }
    mlt_properties_clear(props, "producer");
    mlt_producer_close(producer);
    return NULL;
}

/* Pull a frame as device yuv422; with fix_background_alpha the reference asks for rgba first so
 * that the conversion back creates an alpha plane (:346-367). */
static cuda_image pull_yuv422(mlt_frame frame, int via_rgba, int *w, int *h, int a_role)
{
    if (via_rgba && mlt_frame_has_convert_image(frame)) {
        if (!(a_role ? cuda_frame_get_image_a(frame, mlt_image_rgba, w, h) : cuda_frame_get_image(frame, mlt_image_rgba, w, h, 0)))
            return NULL;
    }
    return a_role ? cuda_frame_get_image_a(frame, mlt_image_yuv422, w, h) : cuda_frame_get_image(frame, mlt_image_yuv422, w, h, 0);
}

static int luma_get_image_impl(mlt_frame a_frame, uint8_t **image, mlt_image_format *format, int *width, int *height,
                          int writable)
{
    mlt_frame b_frame = mlt_frame_pop_frame(a_frame);
    mlt_transition self = mlt_frame_pop_service(a_frame);
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_properties a_props = MLT_FRAME_PROPERTIES(a_frame), b_props = MLT_FRAME_PROPERTIES(b_frame);
    mlt_service service = MLT_TRANSITION_SERVICE(self);
    if (!cuda_thread_ctx())
        return 1;

    mlt_service_lock(service);
    int lw = mlt_properties_get_int(props, "width"), lh = mlt_properties_get_int(props, "height");
    uint16_t *bitmap = mlt_properties_get_data(props, "bitmap", NULL);
    if (lw == 0 || lh == 0)
        lw = *width, lh = *height;
    mlt_producer producer = refresh_map(self, a_frame, &bitmap, &lw, &lh);
    const int have_map = lw > 0 && lh > 0 && bitmap;

    float mix = mlt_transition_get_progress(self, a_frame);
    float frame_delta = mlt_transition_get_progress_delta(self, a_frame);
    float softness = mlt_properties_get_double(props, "softness");
    int progressive = mlt_properties_get_int(a_props, "consumer.progressive") || mlt_properties_get_int(props, "progressive")
                      || mlt_properties_get_int(b_props, "luma.progressive");
    int tff = mlt_properties_get_int(b_props, "top_field_first");
    int reverse = mlt_properties_get_int(props, "reverse");
    int invert = mlt_properties_get_int(props, "invert");
    int alpha_over = mlt_properties_get_int(props, "alpha_over");
    int fix_bg = mlt_properties_get_int(props, "fix_background_alpha");
    if (mix >= 1.0)
        mix -= floor(mix);
    if (mlt_properties_get(props, "fixed"))
        mix = mlt_properties_get_double(props, "fixed");
    if (producer) { /* a clip drives the map: hold the lock while its bitmap is in use */
        invert = !invert;
        mix = 0.5f;
    } else {
        mlt_service_unlock(service);
    }

    int rc = 0;
    mlt_frame dst_frame, src_frame;
    mlt_image_format work = mlt_image_yuv422;
    mltcuda_luma_desc d;
    memset(&d, 0, sizeof(d));
    d.batch = 1, d.translucent = -1; /* decided on the device like is_opaque() (:34-41) */
    d.alpha_over = alpha_over;

    if (have_map) { /* luma wipe (:974-1018) */
        reverse = invert ? !reverse : reverse;
        mix = reverse ? 1 - mix : mix;
        frame_delta *= reverse ? -1.0 : 1.0;
        dst_frame = !invert ? a_frame : b_frame, src_frame = !invert ? b_frame : a_frame;
        if (*format == mlt_image_rgba || *format == mlt_image_rgba64)
            work = *format;
        d.luma_width = lw, d.luma_height = lh;
        d.pos = mix, d.frame_delta = frame_delta, d.softness = softness;
        d.field_order = progressive ? -1 : tff;
        d.invert = invert;
    } else { /* dissolve (:1019-1051) */
        if (mlt_properties_get(a_props, "distort"))
            mlt_properties_set(b_props, "distort", mlt_properties_get(a_props, "distort"));
        if (mlt_color_trc_linear == mlt_image_color_trc_id(mlt_properties_get(a_props, "consumer.mlt_color_trc")))
            mix = (reverse || invert) ? 1.0 - pow(1.0 - mix, 2.0) : pow(mix, 2.0);
        mix = (reverse || invert) ? 1 - mix : mix;
        invert = 0;
        dst_frame = a_frame, src_frame = b_frame;
        if (alpha_over && (*format == mlt_image_rgba || *format == mlt_image_rgba64))
            work = *format;
        d.pos = mix, d.field_order = -1;
    }

    int wd = *width, hd = *height, ws = *width, hs = *height;
    cuda_image dst, src;
    /* Reference quirk: with fix_background_alpha the frames are pulled as rgba and converted back with
     * a bare mlt_frame_convert_image (:361-367,150-160), which leaves the frame's "format" property at
     * rgba -- and mlt_frame_get_alpha() answers NULL for rgba frames (mlt_frame.c:551-564).  The alpha
     * planes the conversion created are therefore invisible to the blend: integer path, planes untouched. */
    int hide_alpha_dst = 0, hide_alpha_src = 0;
    if (work == mlt_image_yuv422) {
        hide_alpha_dst = fix_bg && mlt_frame_has_convert_image(dst_frame);
        hide_alpha_src = fix_bg && mlt_frame_has_convert_image(src_frame);
        if (have_map && mlt_properties_get(MLT_FRAME_PROPERTIES(dst_frame), "distort"))
            mlt_properties_set(MLT_FRAME_PROPERTIES(src_frame), "distort",
                               mlt_properties_get(MLT_FRAME_PROPERTIES(dst_frame), "distort"));
        dst = pull_yuv422(dst_frame, fix_bg, &wd, &hd, 1);
        src = dst ? pull_yuv422(src_frame, fix_bg, &ws, &hs, 0) : NULL;
    } else {
        dst = cuda_frame_get_image_a(dst_frame, work, &wd, &hd);
        src = dst ? cuda_frame_get_image(src_frame, work, &ws, &hs, 0) : NULL;
    }
    if (!dst || !src) {
        rc = 1;
    } else if (dst != src) {
        /* launches go to the context that owns the destination's buffers */
        mltcuda_ctx *ctx = dst->ctx;
        rc = cuda_image_localize(src, ctx);
        if (!rc && have_map && !(d.map = cuda_cached_map(props, "luma.map", ctx, bitmap, lw, lh, mlt_properties_get_int(props, CUDA_MAP_GENERATION))))
            rc = 1;
        d.format = work;
        d.width = *width, d.height = *height;
        d.width_dest = wd, d.height_dest = hd, d.width_src = ws, d.height_src = hs;
        /* Deferred: the opaque progressive yuv422 wipe joins the destination's op
         * list (one fused k_tractor launch per frame).  Opaque is known on the host when neither
         * image has an alpha plane (is_opaque(NULL), :34-41); everything else -- alpha planes (the
         * float path), fields, rgba / rgba64, the dissolves -- flushes the list and launches K3. */
        const int defer = !rc && have_map && work == mlt_image_yuv422 && d.field_order < 0
                          && (hide_alpha_dst || !cuda_image_has_alpha(dst)) && (hide_alpha_src || !src->alpha) && wd == *width
                          && hd == *height && cuda_image_can_defer(dst);
        if (!rc && defer) {
            mltcuda_tractor_op top;
            memset(&top, 0, sizeof(top));
            d.src = src->data, d.translucent = 0;
            top.kind = MLTCUDA_TOP_LUMA, top.u.luma = d;
            rc = cuda_image_defer(dst, &top, src);
        } else if (!rc && !(rc = cuda_image_flush(dst))) {
            d.dest = dst->data, d.src = src->data;
            d.alpha_dest = hide_alpha_dst ? NULL : dst->alpha, d.alpha_src = hide_alpha_src ? NULL : src->alpha;
            CUDA_STAT_COUNT(CS_K_LUMA, 1);
            if (mltcuda_luma(ctx, &d)) {
                mlt_log_error(service, "Unable to dissolve incompatible images (%s)\n", mltcuda_last_error());
                rc = 1;
            }
            cuda_image_used_by(src, ctx);
        }
    }
    if (producer)
        mlt_service_unlock(service);
    if (rc)
        return rc;

    /* the result is the destination frame's image (:1065-1067); when the roles were swapped the
     * handle stays owned by the B frame and A only refers to it */
    if (dst_frame != a_frame)
        mlt_properties_set(a_props, CUDA_IMAGE_TAG, CUDA_IMAGE_TAG_VALUE);
    *image = (uint8_t *) dst, *format = mlt_image_private;
    *width = dst->width, *height = dst->height;
    return 0;
}

static int luma_get_image(mlt_frame a_frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = luma_get_image_impl(a_frame, image, format, width, height, writable);
    CUDA_STAT_END(CS_LUMA, st0);
    return rc;
}

static mlt_frame luma_process(mlt_transition self, mlt_frame a_frame, mlt_frame b_frame)
{
    mlt_frame_push_service(a_frame, self);
    mlt_frame_push_frame(a_frame, b_frame);
    mlt_frame_push_get_image(a_frame, luma_get_image);
    mlt_frame_prepend_convert_image(a_frame, cuda_convert_image);
    mlt_frame_prepend_convert_image(b_frame, cuda_convert_image);
    return a_frame;
}

mlt_transition cuda_transition_luma_init(mlt_profile profile, mlt_service_type type, const char *id, char *lumafile)
{
    cuda_register_get_image(luma_get_image);
    mlt_transition self = mlt_transition_new();
    if (!self)
        return NULL;
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    self->process = luma_process;
    mlt_properties_set(props, "factory", mlt_environment("MLT_PRODUCER"));
    mlt_properties_set(props, "resource", lumafile);
    mlt_properties_set_int(props, "_transition_type", 1);
    return self;
}
