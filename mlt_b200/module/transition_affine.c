/*
 * transition_affine.c (cuda module) -- the `affine` transition on the GPU.
 *
 * Drop-in for src/modules/plus/transition_affine.c of the reference: same service name and
 * properties (rect, fill, distort, keyed, rotate_x/y/z, fix_rotate_*, shear_*, fix_shear_*, ox, oy,
 * scale, scale_x/y, invert_scale, b_alpha, halign, valign, cycle, mirror, mirror_off, repeat_off,
 * b_scaled, always_active; `threads` is accepted and ignored).  The host part evaluates the
 * properties and builds the 3x3 matrix in double with the same operation order as get_affine /
 * affine_* (:93-210, 268-326, 414-714); the pixel work is mltcuda_affine (K4), rgba / rgba64.
 */
#include "cuda_module.h"

#include <ctype.h>
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define AFFINE_MAX_DIMENSION 16000 /* MLT_AFFINE_MAX_DIMENSION (:32) */

typedef double mat3[3][3];

/* a <- a x m^T, the reference's affine_multiply (:108-127) */
static void mat_apply(mat3 a, mat3 m)
{
    mat3 o;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            o[i][j] = a[i][0] * m[j][0] + a[i][1] * m[j][1] + a[i][2] * m[j][2];
    memcpy(a, o, sizeof(o));
}

static void mat_rotate_x(mat3 a, double deg)
{
    double c = cos(deg * M_PI / 180), s = sin(deg * M_PI / 180);
    mat3 m = {{c, 0 - s, 0}, {s, c, 0}, {0, 0, 1}};
    mat_apply(a, m);
}
static void mat_rotate_y(mat3 a, double deg)
{
    double c = cos(deg * M_PI / 180), s = sin(deg * M_PI / 180);
    mat3 m = {{c, 0, 0 - s}, {0, 1, 0}, {s, 0, c}};
    mat_apply(a, m);
}
static void mat_rotate_z(mat3 a, double deg)
{
    double c = cos(deg * M_PI / 180), s = sin(deg * M_PI / 180);
    mat3 m = {{1, 0, 0}, {0, c, s}, {0, -s, c}};
    mat_apply(a, m);
}
static void mat_shear(mat3 a, double x, double y, double z)
{
    mat3 m = {{1, tan(x * M_PI / 180), 0}, {tan(y * M_PI / 180), 1, tan(z * M_PI / 180)}, {0, 0, 1}};
    mat_apply(a, m);
}
static void mat_scale(mat3 a, double sx, double sy)
{
    mat3 m = {{sx, 0, 0}, {0, sy, 0}, {0, 0, 1}};
    mat_apply(a, m);
}
static double map_x(mat3 a, double x, double y)
{
    return a[0][0] * x + a[0][1] * y + a[0][2];
}
static double map_y(mat3 a, double x, double y)
{
    return a[1][0] * x + a[1][1] * y + a[1][2];
}

static double parse_align(const char *s)
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

/* fold `position` into the animation's length with the repeat / mirror rules (:50-72) */
static mlt_position fold_position(mlt_properties props, const char *name, mlt_position position, int length)
{
    mlt_properties_anim_get_double(props, name, position, length); /* parse / refresh */
    mlt_animation anim = mlt_properties_get_animation(props, name);
    if (anim) {
        int alen = mlt_animation_get_length(anim);
        if (!mlt_properties_get_int(props, "repeat_off") && position >= alen && alen != 0) {
            int section = position / alen;
            position -= section * alen;
            if (!mlt_properties_get_int(props, "mirror_off") && section % 2 == 1)
                position = alen - position;
        }
    }
    return position;
}

static double anim_angle(mlt_properties props, const char *name, mlt_position position, mlt_position length)
{
    if (!mlt_properties_get(props, name))
        return 0.0;
    position = fold_position(props, name, position, length);
    double v = mlt_properties_anim_get_double(props, name, position, length);
    return strchr(mlt_properties_get(props, name), '%') ? v * 360 : v;
}

static void build_matrix(mat3 m, mlt_properties props, double position, int length, double scale_w, double scale_h)
{
    double ox, oy;
    if (!mlt_properties_get_int(props, "keyed")) {
        double frx = anim_angle(props, "fix_rotate_x", position, length);
        double fry = anim_angle(props, "fix_rotate_y", position, length);
        double frz = anim_angle(props, "fix_rotate_z", position, length);
        double rx = mlt_properties_get_double(props, "rotate_x");
        double ry = mlt_properties_get_double(props, "rotate_y");
        double rz = mlt_properties_get_double(props, "rotate_z");
        double fsx = anim_angle(props, "fix_shear_x", position, length);
        double fsy = anim_angle(props, "fix_shear_y", position, length);
        double fsz = anim_angle(props, "fix_shear_z", position, length);
        double sx = mlt_properties_get_double(props, "shear_x");
        double sy = mlt_properties_get_double(props, "shear_y");
        double sz = mlt_properties_get_double(props, "shear_z");
        ox = mlt_properties_anim_get_double(props, "ox", position, length);
        oy = mlt_properties_anim_get_double(props, "oy", position, length);
        mat_rotate_x(m, frx + rx * position);
        mat_rotate_y(m, fry + ry * position);
        mat_rotate_z(m, frz + rz * position);
        mat_shear(m, fsx + sx * position, fsy + sy * position, fsz + sz * position);
    } else {
        double rx = anim_angle(props, "rotate_x", position, length);
        double ry = anim_angle(props, "rotate_y", position, length);
        double rz = anim_angle(props, "rotate_z", position, length);
        double sx = anim_angle(props, "shear_x", position, length);
        double sy = anim_angle(props, "shear_y", position, length);
        double sz = anim_angle(props, "shear_z", position, length);
        ox = mlt_properties_anim_get_double(props, "ox", fold_position(props, "ox", position, length), length);
        oy = mlt_properties_anim_get_double(props, "oy", fold_position(props, "oy", position, length), length);
        mat_rotate_x(m, rx);
        mat_rotate_y(m, ry);
        mat_rotate_z(m, rz);
        mat_shear(m, sx, sy, sz);
    }
    m[0][2] += ox * scale_w; /* affine_offset */
    m[1][2] += oy * scale_h;
}

/* affine_max_output (:229-264): extent of the transformed frame, for scale=1 */
static void max_output(mat3 a, double *w, double *h, double dz, double mw, double mh)
{
    int xs[4] = {map_x(a, -mw, mh) / dz, map_x(a, mw, mh) / dz, map_x(a, -mw, -mh) / dz, map_x(a, mw, -mh) / dz};
    int ys[4] = {map_y(a, -mw, mh) / dz, map_y(a, mw, mh) / dz, map_y(a, -mw, -mh) / dz, map_y(a, mw, -mh) / dz};
    int x0 = xs[0], x1 = xs[0], y0 = ys[0], y1 = ys[0];
    for (int i = 1; i < 4; i++) {
        x0 = MIN(x0, xs[i]), x1 = MAX(x1, xs[i]);
        y0 = MIN(y0, ys[i]), y1 = MAX(y1, ys[i]);
    }
    *w = (double) (x1 - x0 + 1) / mw / 2.0;
    *h = (double) (y1 - y0 + 1) / mh / 2.0;
}

static int affine_get_image_impl(mlt_frame a_frame, uint8_t **image, mlt_image_format *format, int *width, int *height,
                            int writable)
{
    mlt_frame b_frame = mlt_frame_pop_frame(a_frame);
    mlt_transition self = mlt_frame_pop_service(a_frame);
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_properties a_props = MLT_FRAME_PROPERTIES(a_frame), b_props = MLT_FRAME_PROPERTIES(b_frame);
    mlt_service service = MLT_TRANSITION_SERVICE(self);
    if (!cuda_thread_ctx())
        return 1;
    const mlt_image_format work = *format == mlt_image_rgba64 ? mlt_image_rgba64 : mlt_image_rgba;

    int b_width = mlt_properties_get_int(b_props, "meta.media.width");
    int b_height = mlt_properties_get_int(b_props, "meta.media.height");
    double b_ar = mlt_frame_get_aspect_ratio(b_frame);
    double b_dar = b_ar * b_width / b_height; /* NaN for producers without meta.media.* (App. B-3) */

    mlt_position position = mlt_transition_get_position(self, a_frame);
    int mirror = mlt_properties_get_position(props, "mirror");
    int length = mlt_transition_get_length(self);
    if (mlt_properties_get_int(props, "always_active")) {
        mlt_properties pp = mlt_properties_get_data(b_props, "_producer", NULL);
        length = mlt_properties_get_int(pp, "out") - mlt_properties_get_int(pp, "in") + 1;
    }
    mlt_profile profile = mlt_service_profile(service);
    const int nw = profile->width, nh = profile->height;
    const double consumer_ar = mlt_profile_sar(profile);
    if (mirror && position > length / 2)
        position = abs(position - length);

    /* preview scaling is disabled while offsets are active (:468-475) */
    double ox = mlt_properties_anim_get_double(props, "ox", position, length);
    double oy = mlt_properties_anim_get_double(props, "oy", position, length);
    if (ox != 0.0 || oy != 0.0)
        *width = nw, *height = nh;

    cuda_image a = cuda_frame_get_image_a(a_frame, work, width, height);
    if (!a)
        return 1;
    *image = (uint8_t *) a, *format = mlt_image_private;

    double scale_w = mlt_profile_scale_width(profile, *width);
    double scale_h = mlt_profile_scale_height(profile, *height);
    mlt_rect r = {0, 0, nw, nh, 1.0};

    mlt_service_lock(service);
    if (mlt_properties_get(props, "rect")) {
        double cycle = mlt_properties_get_double(props, "cycle");
        if (cycle >= 1)
            length = cycle;
        else if (cycle > 0)
            length *= cycle;
        mlt_position ap = fold_position(props, "rect", position, length);
        r = mlt_properties_anim_get_rect(props, "rect", ap, length);
        if (strchr(mlt_properties_get(props, "rect"), '%')) {
            r.x *= nw, r.y *= nh, r.w *= nw, r.h *= nh;
        }
        r.o = (r.o == DBL_MIN) ? 1.0 : MIN(r.o, 1.0);
    }
    /* The reference keeps the service locked through the B fetch and the transform unless
     * threads == 1 (:511-514, 724-727), because its transform occupies the shared slice pool.  The
     * GPU transform does not, and holding the lock over B's upload serialised the worker threads
     * (cfg4 through melt: 190 -> frames/s of one thread).  Unlock here, as the reference does for
     * threads == 1: everything below is per-frame state or internally locked property reads. */
    mlt_service_unlock(service);
    r.x *= scale_w, r.y *= scale_h, r.w *= scale_w, r.h *= scale_h;
    const double geometry_w = r.w, geometry_h = r.h;
    const int fill = mlt_properties_get_int(props, "fill");
    const int distort = mlt_properties_get_int(props, "distort");
    if (!fill) {
        double geometry_dar = r.w * consumer_ar / r.h;
        if (b_dar > geometry_dar) {
            r.w = MIN(r.w, b_width * b_ar / consumer_ar);
            r.h = r.w * consumer_ar / b_dar;
        } else {
            r.h = MIN(r.h, b_height);
            r.w = r.h * b_dar / consumer_ar;
        }
    }

    /* size to request from B (:539-590) */
    if (scale_w != 1.0 || scale_h != 1.0) {
        b_height = CLAMP(*height, 1, AFFINE_MAX_DIMENSION);
        b_width = MAX(b_height * b_dar / b_ar, 1);
        if (b_width > AFFINE_MAX_DIMENSION) {
            b_width = CLAMP(*width, 1, AFFINE_MAX_DIMENSION);
            b_height = MAX(b_width * b_ar / b_dar, 1);
        }
        mlt_properties_set(b_props, "consumer.rescale", mlt_properties_get(a_props, "consumer.rescale"));
        mlt_properties_set_int(b_props, "distort", 1);
    } else if (mlt_properties_get_int(b_props, "always_scale")
               || (!mlt_properties_get_int(b_props, "interpolation_not_required")
                   && (fill || distort || b_width > r.w || b_height > r.h || mlt_properties_get_int(props, "b_scaled")))) {
        b_height = CLAMP(r.h, 1, AFFINE_MAX_DIMENSION);
        b_width = MAX(b_height * b_dar / b_ar, 1);
        if (b_width > AFFINE_MAX_DIMENSION) {
            b_width = CLAMP(r.w, 1, AFFINE_MAX_DIMENSION);
            b_height = MAX(b_width * b_ar / b_dar, 1);
        }
        mlt_properties_set(b_props, "consumer.rescale", mlt_properties_get(a_props, "consumer.rescale"));
        mlt_properties_set_int(b_props, "distort", 1);
    } else {
        mlt_properties_set_int(b_props, "rescale_width", b_width);
        mlt_properties_set_int(b_props, "rescale_height", b_height);
        const char *res = mlt_properties_get(MLT_PRODUCER_PROPERTIES(mlt_frame_get_original_producer(b_frame)), "resource");
        if (res && !strcmp("<track>", res))
            mlt_properties_set(b_props, "consumer.rescale", mlt_properties_get(a_props, "consumer.rescale"));
        else
            mlt_properties_set(b_props, "consumer.rescale", "none");
    }
    mlt_log_debug(service, "requesting image B at resolution %dx%d\n", b_width, b_height);
    mlt_properties_set_int(b_props, "consumer.progressive", 1);

    int rc = 0;
    cuda_image b = cuda_frame_get_image(b_frame, work, &b_width, &b_height, 0);
    if (!b) {
        mlt_frame_set_image(b_frame, NULL, 0, NULL);
        return 1;
    }

    double scale_x = mlt_properties_anim_get_double(props, "scale_x", position, length);
    double scale_y = mlt_properties_anim_get_double(props, "scale_y", position, length);
    const int scale = mlt_properties_get_int(props, "scale");
    const double gsx = (double) b_width / r.w, gsy = (double) b_height / r.h;

    /* launches go to the context that owns A's buffers; B is made readable from it */
    mltcuda_ctx *ctx = a->ctx;
    if (cuda_image_localize(b, ctx)) {
        mlt_frame_set_image(b_frame, NULL, 0, NULL);
        return 1;
    }
    /* Deferred: the transform joins A's op list and runs inside the frame's one
     * fused k_tractor launch.  The fused kernel takes 8-bit rgba and a B image of at least 2x2;
     * anything else (rgba64, width % 8, a degenerate B) flushes A's list and launches K4 here. */
    const int defer = work == mlt_image_rgba && b_width >= 2 && b_height >= 2 && cuda_image_can_defer(a);
    if (!defer && cuda_image_flush(a)) {
        mlt_frame_set_image(b_frame, NULL, 0, NULL);
        return 1;
    }
    mltcuda_affine_desc d;
    memset(&d, 0, sizeof(d));
    d.format = work, d.a_image = defer ? NULL : a->data, d.b_image = b->data, d.batch = 1;
    d.a_width = *width, d.a_height = *height, d.b_width = b_width, d.b_height = b_height;
    d.lower_x = -(r.x + r.w / 2.0), d.lower_y = -(r.y + r.h / 2.0);
    d.mix = r.o;
    d.x_offset = (double) b_width / 2.0, d.y_offset = (double) b_height / 2.0;
    d.b_alpha = mlt_properties_get_int(props, "b_alpha");
    d.minima = 0, d.xmax = b_width - 1, d.ymax = b_height - 1;
    d.interp = MLTCUDA_INTERP_BL;
    if (mlt_properties_get(props, "halign") || mlt_properties_get(props, "valign")) {
        double ha = parse_align(mlt_properties_get(props, "halign"));
        double va = parse_align(mlt_properties_get(props, "valign"));
        d.x_offset = ha * b_width / 2.0, d.y_offset = va * b_height / 2.0;
        d.lower_x = -(r.x + geometry_w * ha / 2.0f), d.lower_y = -(r.y + geometry_h * va / 2.0f);
    }
    mat3 m = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    build_matrix(m, props, (double) position, length, scale_w, scale_h);
    d.dz = m[2][2]; /* MapZ(0,0) */
    if ((int) fabs(d.dz * 1000) < 25)
        goto out; /* edge-on: nothing to draw (:651-656) */

    if (mlt_properties_get_int(props, "invert_scale"))
        scale_x = 1.0 / scale_x, scale_y = 1.0 / scale_y;
    if (distort) {
        scale_x = gsx * (scale_x == 0 ? 1 : scale_x);
        scale_y = gsy * (scale_y == 0 ? 1 : scale_y);
    } else {
        double consumer_dar = consumer_ar * nw / nh;
        if (b_dar > consumer_dar) {
            scale_x = gsx * (scale_x == 0 ? 1 : scale_x);
            scale_y = gsx * (scale_y == 0 ? 1 : scale_y);
            scale_y *= b_ar / consumer_ar;
        } else {
            scale_x = gsy * (scale_x == 0 ? 1 : scale_x);
            scale_y = gsy * (scale_y == 0 ? 1 : scale_y);
            scale_x *= consumer_ar / b_ar;
        }
    }
    if (scale) {
        double sw, sh;
        max_output(m, &sw, &sh, d.dz, *width, *height);
        mat_scale(m, sw * MIN(gsx, gsy), sh * MIN(gsx, gsy));
    } else if (scale_x != 0 && scale_y != 0) {
        mat_scale(m, scale_x, scale_y);
    }
    memcpy(d.matrix, m, sizeof(m));

    const char *interps = mlt_properties_get(a_props, "consumer.rescale");
    if (!interps || !strcmp(interps, "nearest") || !strcmp(interps, "neighbor") || !strcmp(interps, "tiles")
        || !strcmp(interps, "fast_bilinear")) {
        d.interp = MLTCUDA_INTERP_NN; /* lrintf: >= -0.5 and < max + 0.5 */
        d.minima -= 0.5, d.xmax += 0.49, d.ymax += 0.49;
    } else if (!strcmp(interps, "bilinear")) {
        d.interp = MLTCUDA_INTERP_BL;
    } else if (!strcmp(interps, "bicubic") || !strcmp(interps, "hyper") || !strcmp(interps, "sinc")
               || !strcmp(interps, "lanczos") || !strcmp(interps, "spline")) {
        d.interp = MLTCUDA_INTERP_BC; /* ceilf: > -1 and <= max */
        d.minima -= 1;
    }
    if (defer) {
        mltcuda_tractor_op top;
        memset(&top, 0, sizeof(top));
        top.kind = MLTCUDA_TOP_AFFINE, top.u.affine = d;
        rc = cuda_image_defer(a, &top, b); /* the list keeps its own reference to B until it has run */
        goto out;
    }
    CUDA_STAT_COUNT(CS_K_AFFINE, 1);
    if (mltcuda_affine(ctx, &d)) {
        mlt_log_error(service, "affine kernel failed: %s\n", mltcuda_last_error());
        rc = 1;
    }
    cuda_image_used_by(b, ctx);
out:
    /* drop the (potentially large) B image: stream-ordered free, safe after the launch */
    mlt_frame_set_image(b_frame, NULL, 0, NULL);
    mlt_properties_set(b_props, CUDA_IMAGE_TAG, NULL);
    return rc;
}

static int affine_get_image(mlt_frame a_frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = affine_get_image_impl(a_frame, image, format, width, height, writable);
    CUDA_STAT_END(CS_AFFINE, st0);
    return rc;
}

static mlt_frame affine_process(mlt_transition self, mlt_frame a_frame, mlt_frame b_frame)
{
    mlt_frame_push_service(a_frame, self);
    mlt_frame_push_frame(a_frame, b_frame);
    mlt_frame_push_get_image(a_frame, affine_get_image);
    mlt_frame_prepend_convert_image(a_frame, cuda_convert_image);
    mlt_frame_prepend_convert_image(b_frame, cuda_convert_image);
    return a_frame;
}

mlt_transition cuda_transition_affine_init(mlt_profile profile, mlt_service_type type, const char *id, char *arg)
{
    cuda_register_get_image(affine_get_image);
    mlt_transition self = mlt_transition_new();
    if (!self)
        return NULL;
    mlt_properties props = MLT_TRANSITION_PROPERTIES(self);
    mlt_properties_set_int(props, "distort", 0);
    mlt_properties_set(props, "rect", arg ? arg : "0%/0%:100%x100%:100%");
    mlt_properties_set_int(props, "_transition_type", 1);
    mlt_properties_set_int(props, "fill", 1);
    self->process = affine_process;
    return self;
}
