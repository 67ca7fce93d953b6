/*
 * transition_composite.c (cuda module) -- the `composite` transition on the GPU.
 *
 * Drop-in for src/modules/core/transition_composite.c of the reference: same service name, same
 * properties (geometry, halign/valign, aligned, distort, fill, crop, crop_to_fill, titles, luma,
 * luma_invert, softness, operator, alpha_a/alpha_b, progressive, deinterlace, invert, no_alpha,
 * always_active, cycle, mirror_off, repeat_off), same frame-stack protocol.  The host part below
 * evaluates the properties per frame like composite_calculate / get_b_frame_image /
 * transition_get_image (:730-1161) do; the pixel work is mltcuda_composite_setup +
 * mltcuda_composite (K2) on device-resident images.  yuv422 only, like the reference.
 */
#include "cuda_module.h"

#include <ctype.h>
#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct
{
    mlt_rect item;
    int nw, nh;         /* normalised (profile) size */
    int sw, sh;         /* scaled B size before consumer scale */
    int halign, valign; /* 0 left/top, 1 centre, 2 right/bottom */
    int x_src, y_src;   /* crop pan */
} comp_geom;

static int parse_align(const char *s)
{
    if (!s)
        return 0;
    if (isdigit((unsigned char) s[0]))
        return atoi(s);
    if (s[0] == 'c' || s[0] == 'm')
        return 1;
    if (s[0] == 'r' || s[0] == 'b')
        return 2;
    return 0;
}

static void apply_alignment(comp_geom *g)
{
    g->item.x += (g->item.w - g->sw) * g->halign / 2;
    g->item.y += (g->item.h - g->sh) * g->valign / 2;
}

static int prop_or(mlt_properties p, const char *first, const char *second)
{
    int v = mlt_properties_get_int(p, first);
    return v ? v : mlt_properties_get_int(p, second);
}

/* Animated rect with the transition's cycle / repeat / mirror rules (:873-897, :834-855). */
static mlt_rect anim_rect(mlt_transition self, const char *name, double *position)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_position length = mlt_transition_get_length(self);
    double cycle = mlt_properties_get_double(props, "cycle");
    if (cycle >= 1)
        length = cycle;
    else if (cycle > 0)
        length *= cycle;
    /* first evaluation makes mlt_properties parse / refresh the animation */
    mlt_properties_anim_get_rect(props, name, *position, length);
    mlt_animation anim = mlt_properties_get_animation(props, name);
    int alen = mlt_animation_get_length(anim);
    if (!mlt_properties_get_int(props, "repeat_off") && *position >= alen && alen != 0) {
        int section = *position / alen;
        *position -= section * alen;
        if (!mlt_properties_get_int(props, "mirror_off") && section % 2 == 1)
            *position = alen - *position;
    }
    return mlt_properties_anim_get_rect(props, name, *position, length);
}

static void calc_geometry(mlt_transition self, comp_geom *g, double position)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_profile profile = mlt_service_profile(MLT_TRANSITION_SERVICE(self));
    g->nw = profile->width, g->nh = profile->height;
    g->item = anim_rect(self, "geometry", &position);
    const char *text = mlt_properties_get(props, "geometry");
    if (text && strchr(text, '%')) {
        g->item.x *= g->nw, g->item.y *= g->nh;
        g->item.w *= g->nw, g->item.h *= g->nh;
    }
    g->item.o = 100.0 * (g->item.o == DBL_MIN ? 1.0 : MIN(g->item.o, 1.0));
    g->halign = parse_align(mlt_properties_get(props, "halign"));
    g->valign = parse_align(mlt_properties_get(props, "valign"));
    g->x_src = g->y_src = 0;
    if (mlt_properties_get(props, "crop")) { /* :820-857 (position already folded) */
        mlt_rect crop = anim_rect(self, "crop", &position);
        if (strchr(mlt_properties_get(props, "crop"), '%'))
            crop.x *= profile->width, crop.y *= profile->height;
        g->x_src = rint(crop.x), g->y_src = rint(crop.y);
    }
}

/* Decide the size to ask B for (:730-808) and pull it as a device image. */
static cuda_image pull_b(mlt_transition self, mlt_frame b_frame, int *width, int *height, comp_geom *g)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_properties b_props = MLT_FRAME_PROPERTIES(b_frame);
    int resize_alpha = mlt_properties_get_int(b_props, "resize_alpha");
    double out_ar = mlt_profile_sar(mlt_service_profile(MLT_TRANSITION_SERVICE(self)));
    int real_w = prop_or(b_props, "meta.media.width", "width");
    int real_h = prop_or(b_props, "meta.media.height", "height");
    double in_ar = mlt_properties_get_double(b_props, "aspect_ratio");
    int sw = rint((in_ar == 0.0 ? out_ar : in_ar) / out_ar * real_w), sh = real_h;
    int nw = g->item.w, nh = g->item.h;

    if (mlt_properties_get(props, "crop")) {
        g->sw = sw, g->sh = sh; /* no scaling: the rectangle crops B */
    } else if (mlt_properties_get_int(props, "crop_to_fill")) {
        if (sh > 0 && sw * nh / sh >= nw) {
            sw = rint(sw * nh / sh), sh = nh;
        } else if (sw > 0) {
            sh = rint(sh * nw / sw), sw = nw;
        }
        g->sw = sw, g->sh = sh;
    } else if (mlt_properties_get_int(props, "aligned") && !mlt_properties_get_int(props, "distort")
               && !mlt_properties_get_int(b_props, "distort")) {
        if (sw > nw) {
            sh = rint(sh * nw / sw), sw = nw;
        }
        if (sh > nh) {
            sw = rint(sw * nh / sh), sh = nh;
        }
        if (mlt_properties_get_int(props, "fill") && sw > 0 && sh > 0) {
            if (sh < nh && sw * nh / sh <= nw) {
                sw = rint(sw * nh / sh), sh = nh;
            } else if (sw < nw && sh * nw / sw < nh) {
                sh = rint(sh * nw / sw), sw = nw;
            }
        }
        g->sw = sw, g->sh = sh;
    } else {
        g->sw = g->item.w, g->sh = g->item.h;
    }
    if (resize_alpha == 0)
        mlt_properties_set_int(b_props, "distort", mlt_properties_get_int(props, "distort"));
    if (!mlt_properties_get_int(props, "aligned"))
        mlt_properties_set_int(b_props, "resize_alpha", 255);
    if (!mlt_properties_get_int(props, "titles") && !mlt_properties_get(props, "crop"))
        apply_alignment(g);
    *width = rint(g->sw * *width / g->nw);
    *width -= *width % 2;
    *height = rint(g->sh * *height / g->nh);
    cuda_image b = cuda_frame_get_image(b_frame, mlt_image_yuv422, width, height, 1);
    if (!mlt_properties_get(props, "crop"))
        g->sw = *width; /* composite_yuv derives the source stride from sw (:821-827) */
    mlt_properties_set_int(b_props, "resize_alpha", resize_alpha);
    return b;
}

/* ---- luma map: loaded on the host like get_luma (:522-707), scaled on the device ------------- */

typedef struct
{
    mltcuda_ctx *ctx;
    uint16_t *dev;
    const uint16_t *orig;
    int w, h, invert, generation;
} scaled_map;

static void scaled_map_close(void *ptr)
{
    scaled_map *m = ptr;
    if (m) {
        if (m->dev)
            mltcuda_free(m->ctx, m->dev);
        free(m);
    }
}

/* Returns the host original map (cached on the transition), or NULL when no luma is set. */
static uint16_t *host_luma(mlt_transition self, int *lw, int *lh)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_profile profile = mlt_service_profile(MLT_TRANSITION_SERVICE(self));
    char *resource = mlt_properties_get(props, "luma"), *named = resource;
    char path[PATH_MAX];
    if (!resource || !resource[0]) {
        if (mlt_properties_get(props, "_luma")) {
            mlt_properties_set_data(props, "_luma.orig_bitmap", NULL, 0, NULL, NULL);
            mlt_properties_set(props, "_luma", NULL);
        }
        return NULL;
    }
    if (strchr(resource, '%')) { /* "%name" = a stock luma of the profile's lumas dir */
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
    }
    char *old = mlt_properties_get(props, "_luma");
    if (old && old[0] && strcmp(resource, old))
        mlt_properties_set_data(props, "_luma.orig_bitmap", NULL, 0, NULL, NULL);
    uint16_t *orig = mlt_properties_get_data(props, "_luma.orig_bitmap", NULL);
    *lw = mlt_properties_get_int(props, "_luma.orig_width");
    *lh = mlt_properties_get_int(props, "_luma.orig_height");
    if (orig)
        return orig;

    char *ext = strrchr(resource, '.');
    if (ext && !strcmp(ext, ".pgm")) {
        if (mlt_luma_map_from_pgm(resource, &orig, lw, lh)) {
            mlt_luma_map gen = mlt_luma_map_new(named); /* unreadable file: a generated pattern */
            if (profile)
                gen->w = profile->width, gen->h = profile->height;
            orig = mlt_luma_map_render(gen);
            *lw = gen->w, *lh = gen->h;
            free(gen);
        }
    } else {
        mlt_producer producer = mlt_factory_producer(profile, mlt_properties_get(props, "factory"), resource);
        if (producer) {
            mlt_properties pp = MLT_PRODUCER_PROPERTIES(producer);
            mlt_properties_set(pp, "eof", "loop");
            mlt_properties_pass(pp, props, "luma.");
            mlt_frame lf = NULL;
            if (mlt_service_get_frame(MLT_PRODUCER_SERVICE(producer), &lf, 0) == 0) {
                uint8_t *img = NULL;
                mlt_image_format f = mlt_image_yuv422;
                *lw = *lh = 0;
                mlt_properties_set(MLT_FRAME_PROPERTIES(lf), "consumer.rescale", "none");
                mlt_frame_get_image(lf, &img, &f, lw, lh, 0);
                if (img && f == mlt_image_yuv422)
                    mlt_luma_map_from_yuv422(img, &orig, *lw, *lh);
                mlt_frame_close(lf);
            }
            mlt_producer_close(producer);
        }
    }
    if (!orig || *lw <= 0 || *lh <= 0)
        return NULL;
    mlt_properties_set_int(props, CUDA_MAP_GENERATION, mlt_properties_get_int(props, CUDA_MAP_GENERATION) + 1);
    mlt_properties_set_data(props, "_luma.orig_bitmap", orig, *lw * *lh * 2, mlt_pool_release, NULL);
    mlt_properties_set_int(props, "_luma.orig_width", *lw);
    mlt_properties_set_int(props, "_luma.orig_height", *lh);
    mlt_properties_set(props, "_luma", resource);
    return orig;
}

/* Device map scaled to width x height (scale_luma :497-520, invert = XOR 0xFFFF). */
static uint16_t *device_luma(mlt_transition self, mltcuda_ctx *ctx, int width, int height)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    int lw = 0, lh = 0;
    uint16_t *orig = host_luma(self, &lw, &lh);
    if (!orig || width <= 0 || height <= 0)
        return NULL;
    int invert = mlt_properties_get_int(props, "luma_invert");
    char name[64];
    snprintf(name, sizeof(name), "_cuda.luma.%p", (void *) ctx);
    scaled_map *m = mlt_properties_get_data(props, name, NULL);
    const int generation = mlt_properties_get_int(props, CUDA_MAP_GENERATION);
    if (m && m->orig == orig && m->w == width && m->h == height && m->invert == invert && m->generation == generation)
        return m->dev;
    uint16_t *dorig = cuda_cached_map(props, "luma.orig", ctx, orig, lw, lh, mlt_properties_get_int(props, CUDA_MAP_GENERATION));
    if (!dorig)
        return NULL;
    m = calloc(1, sizeof(*m));
    if (!m)
        return NULL;
    m->ctx = ctx, m->orig = orig, m->w = width, m->h = height, m->invert = invert, m->generation = generation;
    m->dev = mltcuda_malloc(ctx, (size_t) width * height * 2);
    if (!m->dev || mltcuda_scale_luma(ctx, m->dev, width, height, dorig, lw, lh, invert * 0xffff)) {
        mlt_log_error(MLT_TRANSITION_SERVICE(self), "luma map scaling failed: %s\n", mltcuda_last_error());
        scaled_map_close(m);
        return NULL;
    }
    mlt_properties_set_data(props, name, m, 0, scaled_map_close, NULL);
    return m->dev;
}

/* ---- get_image ------------------------------------------------------------------------------------ */

static int composite_get_image_impl(mlt_frame a_frame, uint8_t **image, mlt_image_format *format, int *width, int *height,
                               int writable)
{
    mlt_frame b_frame = mlt_frame_pop_frame(a_frame);
    mlt_transition self = mlt_frame_pop_service(a_frame);
    double position = mlt_deque_pop_back_double(MLT_FRAME_IMAGE_STACK(a_frame));
    int out = mlt_frame_pop_service_int(a_frame);
    int in = mlt_frame_pop_service_int(a_frame);
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_service service = MLT_TRANSITION_SERVICE(self);

    if (mlt_properties_get_int(props, "always_active")) {
        mlt_events_block(props, props);
        mlt_properties_set_int(props, "in", in);
        mlt_properties_set_int(props, "out", out);
        mlt_events_unblock(props, props);
    }
    if (mlt_properties_get_int(props, "invert")) {
        mlt_frame t = a_frame;
        a_frame = b_frame, b_frame = t;
    }
    *format = mlt_image_yuv422; /* this compositor is yuv422 only (:980) */

    if (!cuda_thread_ctx())
        return 1;
    cuda_image a = NULL, b = NULL;
    if (!b_frame) {
        a = cuda_frame_get_image_a(a_frame, mlt_image_yuv422, width, height);
        goto done;
    }
    mlt_properties a_props = MLT_FRAME_PROPERTIES(a_frame), b_props = MLT_FRAME_PROPERTIES(b_frame);
    comp_geom g;
    double delta = mlt_transition_get_progress_delta(self, a_frame);
    mlt_position length = mlt_transition_get_length(self);
    mlt_profile profile = mlt_service_profile(service);
    int width_b = *width > 0 ? *width : profile->width;
    int height_b = *height > 0 ? *height : profile->height;

    mlt_service_lock(service);
    calc_geometry(self, &g, position);
    mlt_service_unlock(service);

    if (mlt_properties_get_int(props, "deinterlace")) {
        mlt_properties_set_int(a_props, "consumer.progressive", 1);
        mlt_properties_set_int(b_props, "consumer.progressive", 1);
    }
    /* the reference's "nothing to do" shortcut: B replaces A wholesale (:1021-1028) */
    if (mlt_properties_get_int(props, "no_alpha") && g.item.x == 0 && g.item.y == 0 && g.item.w == *width
        && g.item.h == *height && g.item.o == 100) {
        b = cuda_frame_get_image(b_frame, mlt_image_yuv422, width, height, 1);
        if (!b)
            return 1;
        if (!mlt_frame_is_test_card(a_frame)) {
            /* B keeps owning the handle; A only refers to it, exactly like mlt_frame_replace_image */
            mlt_frame_replace_image(a_frame, (uint8_t *) b, mlt_image_private, *width, *height);
            mlt_properties_set(a_props, CUDA_IMAGE_TAG, CUDA_IMAGE_TAG_VALUE);
        }
        *image = (uint8_t *) b, *format = mlt_image_private;
        return 0;
    }
    if (a_frame == b_frame) {
        double ar = mlt_frame_get_aspect_ratio(b_frame);
        b = pull_b(self, b_frame, &width_b, &height_b, &g);
        mlt_properties_set_double(a_props, "aspect_ratio", ar);
    }
    a = cuda_frame_get_image_a(a_frame, mlt_image_yuv422, width, height);
    if (!a)
        return 1;
    if (g.item.o == 0 || (g.item.w == 0 && g.item.h == 0))
        goto done;

    if (!mlt_properties_get(a_props, "dest_width")) {
        mlt_properties_set_int(a_props, "dest_width", *width);
        mlt_properties_set_int(a_props, "dest_height", *height);
        mlt_properties_set_int(b_props, "dest_width", *width);
        mlt_properties_set_int(b_props, "dest_height", *height);
    } else {
        mlt_properties_set_int(b_props, "dest_width", mlt_properties_get_int(a_props, "dest_width"));
        mlt_properties_set_int(b_props, "dest_height", mlt_properties_get_int(a_props, "dest_height"));
    }
    const int titles = mlt_properties_get_int(props, "titles");
    if (titles) {
        if (!mlt_properties_get(b_props, "consumer.rescale"))
            mlt_properties_set(b_props, "consumer.rescale", "hyper");
        width_b = mlt_properties_get_int(a_props, "dest_width");
        height_b = mlt_properties_get_int(a_props, "dest_height");
    }
    if (a == b) /* compositing a frame onto itself: nothing sensible to do (:1063 *image != image_b) */
        goto done;
    if (!b)
        b = pull_b(self, b_frame, &width_b, &height_b, &g);
    if (!b) {
        /* The reference also leaves A untouched when B yields no image (:1063).  Here "no image" includes a device
         * allocation or upload that failed, which should not vanish silently. */
        mlt_log_warning(service, "[cuda] composite: B image unavailable at position %d, layer skipped (%s)\n",
                        (int) mlt_frame_get_position(a_frame), mltcuda_last_error());
        goto done;
    }
    if (b == a)
        goto done;

    const int progressive = mlt_properties_get_int(a_props, "consumer.progressive")
                            || mlt_properties_get_int(props, "progressive");
    const int tff = mlt_properties_get_int(a_props, "top_field_first");
    const double softness = mlt_properties_get_double(props, "softness");
    /* every launch below goes to the context that owns A's buffers; B is made readable from it */
    mltcuda_ctx *ctx = a->ctx;
    if (cuda_image_localize(b, ctx))
        return 1;
    mlt_service_lock(service);
    uint16_t *luma = device_luma(self, ctx, width_b, height_b);
    mlt_service_unlock(service);
    int op = MLTCUDA_OP_OVER;
    const char *opname = mlt_properties_get(props, "operator");
    if (opname) {
        if (!strcmp(opname, "or"))
            op = MLTCUDA_OP_OR;
        if (!strcmp(opname, "and"))
            op = MLTCUDA_OP_AND;
        if (!strcmp(opname, "xor"))
            op = MLTCUDA_OP_XOR;
    }
    /* alpha plane overrides (:1098-1102): they rewrite A's plane, so A's pending ops run first */
    if (mlt_properties_get(props, "alpha_a") && cuda_image_has_alpha(a)) {
        if (cuda_image_flush(a))
            return 1;
        mltcuda_memset(ctx, a->alpha, mlt_properties_get_int(props, "alpha_a"), (size_t) *width * *height);
    }
    if (mlt_properties_get(props, "alpha_b") && b->alpha)
        mltcuda_memset(ctx, b->alpha, mlt_properties_get_int(props, "alpha_b"), (size_t) width_b * height_b);
    /* Deferred: the blend joins A's op list and runs inside the one fused k_tractor
     * launch of the frame.  What the fused kernel refuses (width % 8, see cuda_image_can_defer) runs
     * here as its own K2 launch after A's list has been flushed. */
    const int defer = cuda_image_can_defer(a);
    if (!defer && cuda_image_flush(a))
        return 1;

    for (int field = 0; field < (progressive ? 1 : 2); field++) {
        double field_position = position + field * delta * length;
        int field_id = progressive ? -1 : (tff ? 1 - field : field);
        mlt_service_lock(service);
        calc_geometry(self, &g, field_position);
        mlt_service_unlock(service);
        /* calc_geometry does not know the scaled B size: keep what pull_b negotiated */
        g.sw = mlt_properties_get(props, "crop") ? g.sw : b->width;
        if (titles) {
            g.item.w = rint(*width * (g.item.w / g.nw));
            g.nw = g.item.w;
            g.item.h = rint(*height * (g.item.h / g.nh));
            g.nh = *height;
            g.sw = width_b, g.sh = height_b;
        }
        if (mlt_properties_get(props, "crop")) {
            if (g.x_src == 0)
                width_b = width_b > g.item.w ? g.item.w : width_b;
            if (g.y_src == 0)
                height_b = height_b > g.item.h ? g.item.h : height_b;
        } else if (mlt_properties_get_int(props, "crop_to_fill")) {
            if (g.item.w < g.sw)
                g.x_src = rint((g.item.w - g.sw) * g.halign / 2);
            if (g.item.h < g.sh)
                g.y_src = rint((g.item.h - g.sh) * g.valign / 2);
            if (g.x_src == 0)
                width_b = width_b > g.item.w ? g.item.w : width_b;
            if (g.y_src == 0)
                height_b = height_b > g.item.h ? g.item.h : height_b;
        } else {
            apply_alignment(&g);
        }
        mltcuda_composite_geometry cg = {g.item.x, g.item.y, g.item.w, g.item.h, g.item.o, g.nw, g.nh,
                                         g.sw, g.sh, g.x_src, g.y_src};
        mltcuda_composite_desc d;
        memset(&d, 0, sizeof(d));
        d.op = op, d.batch = 1;
        if (mltcuda_composite_setup(&d, defer ? CUDA_LAZY_BASE : a->data, *width, *height, b->data, width_b, height_b, b->alpha,
                                    defer ? NULL : a->alpha, &cg, field_id, luma, softness))
            continue; /* nothing visible */
        if (defer) {
            mltcuda_tractor_op top;
            memset(&top, 0, sizeof(top));
            top.kind = MLTCUDA_TOP_COMPOSITE, top.u.composite = d;
            if (cuda_image_defer(a, &top, b))
                return 1;
            continue;
        }
        CUDA_STAT_COUNT(CS_K_COMPOSITE, 1);
        if (mltcuda_composite(ctx, &d)) {
            mlt_log_error(service, "composite kernel failed: %s\n", mltcuda_last_error());
            return 1;
        }
    }
    if (!defer)
        cuda_image_used_by(b, ctx);
done:
    if (!a)
        return 1;
    *image = (uint8_t *) a, *format = mlt_image_private;
    *width = a->width, *height = a->height;
    return 0;
}

static int composite_get_image(mlt_frame a_frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = composite_get_image_impl(a_frame, image, format, width, height, writable);
    CUDA_STAT_END(CS_COMPOSITE, st0);
    return rc;
}

static mlt_frame composite_process(mlt_transition self, mlt_frame a_frame, mlt_frame b_frame)
{
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    if (!mlt_properties_get_int(props, "always_active")) {
        mlt_frame_push_service_int(a_frame, mlt_properties_get_int(props, "in"));
        mlt_frame_push_service_int(a_frame, mlt_properties_get_int(props, "out"));
        mlt_deque_push_back_double(MLT_FRAME_IMAGE_STACK(a_frame),
                                   mlt_frame_get_position(a_frame) - mlt_transition_get_in(self));
    } else {
        mlt_properties pp = mlt_properties_get_data(MLT_FRAME_PROPERTIES(b_frame), "_producer", NULL);
        mlt_frame_push_service_int(a_frame, mlt_properties_get_int(pp, "in"));
        mlt_frame_push_service_int(a_frame, mlt_properties_get_int(pp, "out"));
        mlt_deque_push_back_double(MLT_FRAME_IMAGE_STACK(a_frame),
                                   mlt_properties_get_int(pp, "_frame") - mlt_properties_get_int(pp, "in"));
    }
    mlt_frame_push_service(a_frame, self);
    mlt_frame_push_frame(a_frame, b_frame);
    mlt_frame_push_get_image(a_frame, composite_get_image);
    /* make sure the consumer's request for a host format reaches our converter first */
    mlt_frame_prepend_convert_image(a_frame, cuda_convert_image);
    mlt_frame_prepend_convert_image(b_frame, cuda_convert_image);
    return a_frame;
}

mlt_transition cuda_transition_composite_init(mlt_profile profile, mlt_service_type type, const char *id, char *arg)
{
    cuda_register_get_image(composite_get_image);
    mlt_transition self = mlt_transition_new();
    if (!self)
        return NULL;
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    self->process = composite_process;
    mlt_properties_set(props, "geometry", arg ? arg : "0/0:100%x100%");
    mlt_properties_set(props, "factory", mlt_environment("MLT_PRODUCER"));
    mlt_properties_set_int(props, "aligned", 1);
    mlt_properties_set_int(props, "progressive", 1);
    mlt_properties_set_int(props, "_transition_type", 1); /* video only */
    return self;
}
