/*
 * producer_colour.c (cuda module) -- the `colour` / `color` producer with a device-resident picture.
 *
 * Replaces src/modules/core/producer_colour.c:33-345 of the reference where it matters on the
 * compositing path: when one of the CUDA services pulls the frame, the picture is
 * filled on the GPU by mltcuda_colour() (k_colour_*: the reference's constant fill, RGB2YUV_601 for
 * yuv422, (c << 8) + c for rgba64, alpha plane = memset(alpha)) in the format that service works in,
 * and never exists in host memory.  Without this, a UHD rgba colour track costs a 33 MB memcpy in the
 * stock producer plus a 33 MB upload per frame.
 *
 * Every other request -- a host consumer, a CPU filter, the checkerboard pattern, a forced
 * mlt_image_format -- is answered by the stock producer itself: this producer owns a stock `colour`
 * instance, forwards its properties, pulls the frame from it and adopts the image, so behaviour
 * outside the fast path is the reference's by construction.
 */
#include "cuda_module.h"

#include <stdlib.h>
#include <string.h>

/* set while this module creates its inner stock producer, so the factory hook leaves it alone */
__thread int cuda_creating_stock_producer = 0;

static mlt_producer stock_child(mlt_producer self)
{
    mlt_properties props = MLT_PRODUCER_PROPERTIES(self);
    mlt_producer child = mlt_properties_get_data(props, "_cuda_stock", NULL);
    if (!child) {
        cuda_creating_stock_producer++;
        child = mlt_factory_producer(mlt_service_profile(MLT_PRODUCER_SERVICE(self)), "colour",
                                     mlt_properties_get(props, "resource"));
        cuda_creating_stock_producer--;
        if (child)
            mlt_properties_set_data(props, "_cuda_stock", child, 0, (mlt_destructor) mlt_producer_close, NULL);
    }
    return child;
}

static void copy_prop(mlt_properties dst, mlt_properties src, const char *name)
{
    if (mlt_properties_exists(src, name))
        mlt_properties_set(dst, name, mlt_properties_get(src, name));
}

static int colour_get_image_impl(mlt_frame frame, uint8_t **buffer, mlt_image_format *format, int *width, int *height,
                            int writable)
{
    mlt_properties fp = MLT_FRAME_PROPERTIES(frame);
    mlt_producer self = mlt_frame_pop_service(frame);
    mlt_properties props = MLT_PRODUCER_PROPERTIES(self);
    const char *now = mlt_properties_get(props, "resource");
    const int resident = cuda_request_resident(frame, *format);
    /* asked for a resident image: fill in the pulling service's working format (what the stock
     * producer would be asked for).  Through the loader's normalisers the puller is the device
     * `rescale` (filter_rescale.c of this module), which asks for yuv422 as the stock one does
     * (:221-223); the transition then converts on the device exactly as the reference's imageconvert
     * would on the host. */
    const int want = *format == mlt_image_private ? mlt_properties_get_int(fp, CUDA_WORK_FORMAT) : (int) *format;

    if (resident && now && strcmp(now, "checkerboard") && !mlt_properties_get(props, "mlt_image_format")
        && (want == mlt_image_yuv422 || want == mlt_image_rgba || want == mlt_image_rgba64)) {
        mltcuda_ctx *ctx = cuda_thread_ctx();
        if (!ctx)
            return 1;
        if (*width <= 0)
            *width = mlt_service_profile(MLT_PRODUCER_SERVICE(self))->width;
        if (*height <= 0)
            *height = mlt_service_profile(MLT_PRODUCER_SERVICE(self))->height;
        /* Below the loader's rescale the request is for meta.media.width x height (the profile's
         * size, filter_rescale.c:200-204) and the filter then scales that to rescale_width x
         * rescale_height (nearest neighbour; :206-209).  A uniform picture scaled is the same uniform
         * picture: fill at the final size, the rescaler then finds nothing to scale. */
        if (mlt_properties_get_int(fp, "rescale_width") > 0 && mlt_properties_get_int(fp, "rescale_height") > 0) {
            *width = mlt_properties_get_int(fp, "rescale_width");
            *height = mlt_properties_get_int(fp, "rescale_height");
        }
        mlt_color c = mlt_properties_get_color(props, "resource");
        /* alpha plane exactly when the reference makes one for this format (:217-224): never merged
         * into rgba / rgba64 by the services, present for yuv422 only if the colour is translucent */
        cuda_image ci = cuda_image_new(ctx, want, *width, *height, want == mlt_image_yuv422 && c.a < 255);
        if (!ci)
            return 1;
        const int abi_fmt = want == mlt_image_yuv422 ? MLTCUDA_FMT_YUV422 : want == mlt_image_rgba ? MLTCUDA_FMT_RGBA : MLTCUDA_FMT_RGBA64;
        if (mltcuda_colour(ctx, ci->data, ci->alpha, abi_fmt, *width, *height, c.r, c.g, c.b, c.a)
            || cuda_frame_set_image(frame, ci)) {
            mlt_log_error(NULL, "[cuda] colour kernel failed: %s\n", mltcuda_last_error());
            cuda_image_close(ci);
            return 1;
        }
        *buffer = (uint8_t *) ci, *format = mlt_image_private;
        mlt_properties_set_double(fp, "aspect_ratio", mlt_properties_get_double(props, "aspect_ratio"));
        mlt_properties_set_int(fp, "meta.media.width", *width);
        mlt_properties_set_int(fp, "meta.media.height", *height);
        const int yuv = want == mlt_image_yuv422; /* :240-268 */
        mlt_properties_set_int(fp, "colorspace", yuv ? mlt_colorspace_bt601 : mlt_colorspace_rgb);
        mlt_properties_set_int(fp, "color_trc", yuv ? mlt_color_trc_smpte170m : mlt_color_trc_iec61966_2_1);
        mlt_properties_set_int(fp, "color_primaries", yuv ? mlt_color_pri_smpte170m : mlt_color_pri_bt709);
        mlt_properties_set_int(fp, "full_range", yuv ? 0 : 1);
        return 0;
    }

    /* anything else: the stock producer answers, this frame adopts its picture */
    mlt_service_lock(MLT_PRODUCER_SERVICE(self));
    mlt_producer child = stock_child(self);
    mlt_frame cf = NULL;
    if (child) {
        mlt_properties cp = MLT_PRODUCER_PROPERTIES(child);
        copy_prop(cp, props, "resource");
        copy_prop(cp, props, "mlt_image_format");
        mlt_properties_set_double(cp, "aspect_ratio", mlt_properties_get_double(props, "aspect_ratio"));
        mlt_producer_seek(child, mlt_frame_get_position(frame));
        mlt_service_get_frame(MLT_PRODUCER_SERVICE(child), &cf, 0);
    }
    mlt_service_unlock(MLT_PRODUCER_SERVICE(self));
    if (!cf)
        return 1;
    mlt_image_format f = *format == mlt_image_private ? (want ? (mlt_image_format) want : mlt_image_rgba) : *format;
    int err = mlt_frame_get_image(cf, buffer, &f, width, height, writable);
    if (err || !*buffer) {
        mlt_frame_close(cf);
        return 1;
    }
    *format = f;
    mlt_properties cfp = MLT_FRAME_PROPERTIES(cf);
    int size = 0;
    mlt_properties_get_data(cfp, "image", &size);
    mlt_properties_set_data(fp, "_cuda_stock_frame", cf, 0, (mlt_destructor) mlt_frame_close, NULL);
    mlt_frame_set_image(frame, *buffer, size, NULL); /* owned by the stock frame kept above */
    int asize = 0;
    uint8_t *alpha = mlt_properties_get_data(cfp, "alpha", &asize);
    mlt_frame_set_alpha(frame, alpha, asize, NULL);
    static const char *carry[] = {"aspect_ratio", "meta.media.width", "meta.media.height", "colorspace", "color_trc",
                                  "color_primaries", "full_range", NULL};
    for (int i = 0; carry[i]; i++)
        copy_prop(fp, cfp, carry[i]);
    return 0;
}

static int colour_get_image(mlt_frame frame, uint8_t **buffer, mlt_image_format *format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = colour_get_image_impl(frame, buffer, format, width, height, writable);
    CUDA_STAT_END(CS_COLOUR, st0);
    return rc;
}

static int colour_get_frame(mlt_producer producer, mlt_frame_ptr frame, int index)
{
    *frame = mlt_frame_init(MLT_PRODUCER_SERVICE(producer));
    if (*frame) {
        mlt_properties fp = MLT_FRAME_PROPERTIES(*frame);
        mlt_properties props = MLT_PRODUCER_PROPERTIES(producer);
        mlt_profile profile = mlt_service_profile(MLT_PRODUCER_SERVICE(producer));
        mlt_frame_set_position(*frame, mlt_producer_position(producer));
        mlt_properties_set_int(fp, "progressive", 1);
        mlt_properties_set_double(fp, "aspect_ratio", mlt_profile_sar(profile));
        mlt_properties_set_int(fp, "meta.media.width", profile->width);
        mlt_properties_set_int(fp, "meta.media.height", profile->height);
        if (mlt_properties_get(props, "colour")) /* alias (:296-300) */
            mlt_properties_set(props, "resource", mlt_properties_get(props, "colour"));
        char *cs = mlt_properties_get(props, "resource");
        if (cs && strchr(cs, '/')) { /* a path-like resource keeps its last component (:302-307) */
            cs = strdup(strrchr(cs, '/') + 1);
            mlt_properties_set(props, "resource", cs);
            free(cs);
        }
        if (mlt_properties_exists(props, "mlt_image_format")) {
            mlt_properties_set_int(fp, "format", mlt_image_format_id(mlt_properties_get(props, "mlt_image_format")));
        } else {
            mlt_color c = mlt_properties_get_color(props, "resource");
            mlt_properties_set_int(fp, "format", c.a < 255 ? mlt_image_rgba : mlt_image_yuv422);
        }
        mlt_frame_push_service(*frame, producer);
        mlt_frame_push_get_image(*frame, colour_get_image);
        mlt_properties_set_int(fp, "interpolation_not_required", 1);
    }
    mlt_producer_prepare_next(producer);
    return 0;
}

static void colour_close(mlt_producer producer)
{
    producer->close = NULL;
    mlt_producer_close(producer);
    free(producer);
}

mlt_producer cuda_producer_colour_init(mlt_profile profile, mlt_service_type type, const char *id, char *colour)
{
    mlt_producer producer = calloc(1, sizeof(struct mlt_producer_s));
    if (producer && mlt_producer_init(producer, NULL) == 0) {
        mlt_properties props = MLT_PRODUCER_PROPERTIES(producer);
        producer->get_frame = colour_get_frame;
        producer->close = (mlt_destructor) colour_close;
        mlt_properties_set(props, "resource", (!colour || !*colour) ? "0x000000ff" : colour);
        mlt_properties_set(props, "_resource", "");
        mlt_properties_set_double(props, "aspect_ratio", mlt_profile_sar(profile));
        mlt_properties_set_int(props, "_cuda_producer", 1);
        return producer;
    }
    free(producer);
    return NULL;
}
