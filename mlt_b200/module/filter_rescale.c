/* filter_rescale.c (cuda module) -- the loader's `rescale` and `resize` normalisers on the device
 * (SURVEY 8(f) f2).
 *
 * The loader attaches rescale + resize (+ crop, fieldorder, deinterlace) to every producer
 * (src/modules/core/loader.ini, producer_loader.c:232-248), so they sit between every producer and
 * every transition.  The stock filters work on host memory only: rescale overwrites the requested
 * format with yuv422 before pulling its input (src/modules/core/filter_rescale.c:221-223) and scales
 * with filter_scale (:48-120); resize pads whatever comes back smaller than it was asked for
 * (filter_resize.c:143-193) -- handed a device-resident mlt_image_private picture it would "pad" the
 * handle itself.
 *
 * Both stock filters are left in place and keep serving every host request unchanged.  This file
 * wraps their process(): right above the entries the stock process pushes on the frame's image stack
 * it pushes a get_image of its own.  When the request comes from one of this module's services
 * (format mlt_image_private, "_cuda_requested" raised) that get_image discards the stock entries and
 * does the filter's job on the device:
 *   rescale: pulls its input as a device yuv422 image (as the stock filter forces yuv422) and scales
 *            it with mltcuda_rescale_yuv422 / mltcuda_rescale_alpha (K6) when the sizes differ;
 *   resize:  the aspect / letterbox arithmetic of filter_resize.c:196-298 on the host, the pad with
 *            mltcuda_resize_pad in the format the pulling service works in.
 * Anything else falls through to the stock get_image below.
 */
#include "cuda_module.h"

#include <math.h>
#include <string.h>

typedef mlt_frame (*process_fn)(mlt_filter, mlt_frame);

static int is_cuda_request(mlt_frame frame, mlt_image_format format)
{
    return format == mlt_image_private && mlt_properties_get_int(MLT_FRAME_PROPERTIES(frame), CUDA_REQUESTED) > 0;
}

/* drop the n entries the stock process() pushed (get_image, service, ...) */
static void discard_stock_entries(mlt_frame frame, int n)
{
    while (n-- > 0)
        mlt_deque_pop_back(MLT_FRAME_IMAGE_STACK(frame));
}

/* Run the stock process and push our get_image above whatever it pushed. */
static mlt_frame wrap_process(mlt_filter filter, mlt_frame frame, mlt_get_image ours, int with_aspect)
{
    process_fn stock = (process_fn) mlt_properties_get_data(MLT_FILTER_PROPERTIES(filter), "_cuda_stock_process", NULL);
    mlt_deque stack = MLT_FRAME_IMAGE_STACK(frame);
    const double aspect = mlt_frame_get_aspect_ratio(frame); /* what the stock resize records (:306) */
    const int before = mlt_deque_count(stack);
    if (stock)
        frame = stock(filter, frame);
    const int pushed = mlt_deque_count(stack) - before;
    if (with_aspect)
        mlt_deque_push_back_double(stack, aspect);
    mlt_frame_push_service_int(frame, pushed);
    mlt_frame_push_service(frame, filter);
    mlt_frame_push_get_image(frame, ours);
    return frame;
}

/* ---- rescale ------------------------------------------------------------------------------------- */

/* Has somebody replaced the scaler function ("method") core's filter_rescale_init installed
 * (src/modules/core/filter_rescale.c:303-304) since cuda_rescale_install() looked? */
#define CUDA_CORE_SCALER "_cuda_core_method"
static int foreign_scaler(mlt_properties fp)
{
    return mlt_properties_get_data(fp, "method", NULL) != mlt_properties_get_data(fp, CUDA_CORE_SCALER, NULL);
}

static int rescale_get_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    mlt_filter filter = mlt_frame_pop_service(frame);
    const int pushed = mlt_frame_pop_service_int(frame);
    if (!is_cuda_request(frame, *format))
        return mlt_frame_get_image(frame, image, format, width, height, writable); /* the stock rescale */
    mlt_properties fp = MLT_FILTER_PROPERTIES(filter);
    if (foreign_scaler(fp)) {
        /* Not core `rescale` any more: swscale and gtkrescale build themselves on it and then swap the "method"
         * function for their own (avformat/filter_swscale.c:391-402, gdk/filter_rescale.c:140-151), and the
         * loader prefers them (loader.ini: rescaler=movit.resample,swscale,gtkrescale,rescale).  K6 is core's
         * nearest-neighbour scaler only, so their interpolation must stay theirs: run the stock stack for a host
         * picture in the format the pulling service works in; whoever asked for residency uploads the result. */
        mlt_image_format host = (mlt_image_format) mlt_properties_get_int(p, CUDA_WORK_FORMAT);
        if (!cuda_format_supported(host))
            host = mlt_image_yuv422;
        const int requested = mlt_properties_get_int(p, CUDA_REQUESTED);
        mlt_properties_set_int(p, CUDA_REQUESTED, 0);
        const int rc = mlt_frame_get_image(frame, image, &host, width, height, writable);
        mlt_properties_set_int(p, CUDA_REQUESTED, requested);
        *format = host;
        return rc;
    }
    discard_stock_entries(frame, pushed);

    if (*width == 0 || *height == 0) {
        mlt_profile profile = mlt_service_profile(MLT_FILTER_SERVICE(filter));
        *width = profile->width, *height = profile->height;
    }
    if (*width < 6 || *height < 6)
        return 1; /* the stock filter refuses small pictures too (:186,262-264) */
    int iw = *width, ih = *height, ow = *width, oh = *height;
    if (mlt_properties_get(fp, "factor")) {
        const double factor = mlt_properties_get_double(fp, "factor");
        ow *= factor, oh *= factor;
    }
    const char *interps = mlt_properties_get(p, "consumer.rescale");
    if (!interps) {
        interps = mlt_properties_get(fp, "interpolation");
        mlt_properties_set(p, "consumer.rescale", interps);
    }
    if (mlt_properties_get_int(p, "meta.media.width")) {
        iw = mlt_properties_get_int(p, "meta.media.width");
        ih = mlt_properties_get_int(p, "meta.media.height");
    }
    /* what the producer is really being asked for */
    const int scaling = interps && strcmp(interps, "none");
    mlt_properties_set_int(p, "rescale_width", scaling ? *width : iw);
    mlt_properties_set_int(p, "rescale_height", scaling ? *height : ih);
    if (ih != oh && (!interps || strcmp(interps, "nearest") || ih % oh != 0))
        mlt_properties_set_int(p, "consumer.progressive", 1);

    /* the built-in scaler works in yuv422 (:221-223): the input comes up as a device yuv422 image */
    cuda_image ci = cuda_frame_get_image(frame, mlt_image_yuv422, &iw, &ih, writable);
    if (!ci)
        return 1;
    interps = mlt_properties_get(p, "consumer.rescale"); /* the producer may have switched scaling off */
    if (interps && strcmp(interps, "none") && (iw != ow || ih != oh)) {
        ow -= ow % 2; /* pairs of pixels; an odd request never reaches here from resize (:279-282) */
        cuda_image out = cuda_image_new(ci->ctx, mlt_image_yuv422, ow, oh, ci->alpha != NULL);
        int rc = !out;
        /* (filter_scale never writes the last row of an odd-height output: pool memory in the reference,
         * device pool memory here) */
        if (!rc)
            rc = mltcuda_rescale_yuv422(ci->ctx, out->data, ow, oh, ci->data, iw, ih, 1, 0, 0);
        if (!rc && ci->alpha)
            rc = mltcuda_rescale_alpha(ci->ctx, out->alpha, ow, oh, ci->alpha, iw, ih);
        CUDA_STAT_COUNT(CS_K_RESCALE, 1);
        if (rc || cuda_frame_set_image(frame, out)) { /* takes over from ci */
            mlt_log_error(MLT_FILTER_SERVICE(filter), "[cuda] rescale %dx%d -> %dx%d failed: %s\n", iw, ih, ow, oh,
                          mltcuda_last_error());
            cuda_image_close(out);
            return 1;
        }
        ci = out;
    }
    *width = ci->width, *height = ci->height;
    *image = (uint8_t *) ci, *format = mlt_image_private;
    return 0;
}

static mlt_frame rescale_process(mlt_filter filter, mlt_frame frame)
{
    return wrap_process(filter, frame, rescale_get_image, 0);
}

/* ---- resize -------------------------------------------------------------------------------------- */

static int resize_get_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    mlt_filter filter = mlt_frame_pop_service(frame);
    const int pushed = mlt_frame_pop_service_int(frame);
    double aspect_ratio = mlt_deque_pop_back_double(MLT_FRAME_IMAGE_STACK(frame));
    if (!is_cuda_request(frame, *format))
        return mlt_frame_get_image(frame, image, format, width, height, writable); /* the stock resize */
    discard_stock_entries(frame, pushed);

    mlt_profile profile = mlt_service_profile(MLT_FILTER_SERVICE(filter));
    const double consumer_aspect = mlt_profile_sar(profile);
    mlt_image_format work = (mlt_image_format) mlt_properties_get_int(p, CUDA_WORK_FORMAT);
    if (work != mlt_image_rgba && work != mlt_image_rgba64 && work != mlt_image_rgb)
        work = mlt_image_yuv422;
    if (*width == 0 || *height == 0)
        *width = profile->width, *height = profile->height;
    int ow = *width, oh = *height;
    if (aspect_ratio == 0.0)
        aspect_ratio = consumer_aspect;
    mlt_properties_set_double(p, "aspect_ratio", aspect_ratio);

    const char *rescale = mlt_properties_get(p, "consumer.rescale");
    const int no_scaling = rescale && !strcmp(rescale, "none");
    if (!no_scaling) {
        if (mlt_properties_get_int(p, "distort") == 0 && profile) {
            /* fit the source's display aspect into the requested frame (:236-268) */
            int real_w = mlt_properties_get_int(p, "meta.media.width"), real_h = mlt_properties_get_int(p, "meta.media.height");
            if (real_w == 0)
                real_w = mlt_properties_get_int(p, "width");
            if (real_h == 0)
                real_h = mlt_properties_get_int(p, "height");
            const double input_ar = aspect_ratio * real_w / real_h;
            const double output_ar = consumer_aspect * ow / oh;
            int sw = rint((input_ar * profile->width) / output_ar), sh = profile->height;
            if (sw > profile->width) {
                sw = profile->width;
                sh = rint((output_ar * profile->height) / input_ar);
            }
            ow = rint(sw * ow / profile->width);
            oh = rint(sh * oh / profile->height);
            mlt_frame_set_aspect_ratio(frame, consumer_aspect);
        }
        mlt_properties_set_int(p, "distort", 0);
        mlt_properties_set_int(p, "resize_width", *width);
        mlt_properties_set_int(p, "resize_height", *height);
        if (work == mlt_image_yuv422) {
            ow -= ow % 2;
            *width -= *width % 2;
        }
    }
    cuda_image ci = cuda_frame_get_image(frame, work, &ow, &oh, writable);
    if (!ci)
        return 1;
    if (!no_scaling && (ci->width < *width || ci->height < *height) && ci->width <= *width && ci->height <= *height) {
        const int has_alpha = ci->alpha != NULL && work != mlt_image_rgba; /* :176 */
        const int alpha_value = mlt_properties_get_int(p, "resize_alpha") & 0xff;
        cuda_image out = cuda_image_new(ci->ctx, work, *width, *height, has_alpha);
        int rc = !out;
        if (!rc)
            rc = mltcuda_resize_pad(ci->ctx, out->data, *width, *height, ci->data, ci->width, ci->height, work, alpha_value);
        if (!rc && has_alpha)
            rc = mltcuda_resize_pad(ci->ctx, out->alpha, *width, *height, ci->alpha, ci->width, ci->height, MLTCUDA_FMT_NONE,
                                    alpha_value);
        CUDA_STAT_COUNT(CS_K_PAD, 1);
        if (rc || cuda_frame_set_image(frame, out)) {
            mlt_log_error(MLT_FILTER_SERVICE(filter), "[cuda] resize pad %dx%d -> %dx%d failed: %s\n", ci->width, ci->height,
                          *width, *height, mltcuda_last_error());
            cuda_image_close(out);
            return 1;
        }
        ci = out;
    }
    *width = ci->width, *height = ci->height;
    *image = (uint8_t *) ci, *format = mlt_image_private;
    return 0;
}

static mlt_frame resize_process(mlt_filter filter, mlt_frame frame)
{
    return wrap_process(filter, frame, resize_get_image, 1);
}

/* ---- mirror (flip) ----------------------------------------------------------------------------------- */

/* core `mirror` (src/modules/core/filter_mirror.c) works in place on a host yuv422 picture: between a device
 * producer and a CUDA transition it costs a download, the CPU shuffle and an upload per frame (the track-1
 * decorrelator of the benchmark graphs).  Its "flip" mode on an even-width picture is done on the
 * device when a CUDA service is pulling; every other mode, odd widths and every host request stay with the
 * stock code. */
static int mirror_get_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    mlt_filter filter = mlt_frame_pop_service(frame);
    const int pushed = mlt_frame_pop_service_int(frame);
    const char *mode = mlt_properties_get(MLT_FILTER_PROPERTIES(filter), "mirror");
    if (!is_cuda_request(frame, *format) || !mode || strcmp(mode, "flip") || *width % 2)
        return mlt_frame_get_image(frame, image, format, width, height, writable); /* the stock mirror */
    discard_stock_entries(frame, pushed);
    cuda_image ci = cuda_frame_get_image(frame, mlt_image_yuv422, width, height, 1); /* the filter works in yuv422 (:300) */
    if (!ci)
        return 1;
    CUDA_STAT_COUNT(CS_K_MIRROR, 1);
    cuda_image out = cuda_image_new(ci->ctx, mlt_image_yuv422, ci->width, ci->height, ci->alpha != NULL);
    if (!out || mltcuda_mirror_flip(ci->ctx, out->data, out->alpha, ci->data, ci->alpha, ci->width, ci->height)
        || cuda_frame_set_image(frame, out)) {
        mlt_log_error(MLT_FILTER_SERVICE(filter), "[cuda] mirror flip %dx%d failed: %s\n", ci->width, ci->height, mltcuda_last_error());
        cuda_image_close(out);
        return 1;
    }
    ci = out;
    *width = ci->width, *height = ci->height;
    *image = (uint8_t *) ci, *format = mlt_image_private;
    return 0;
}

static mlt_frame mirror_process(mlt_filter filter, mlt_frame frame)
{
    return wrap_process(filter, frame, mirror_get_image, 0);
}

/* ---- installation (factory.c: "filter-create-done") ------------------------------------------------- */

static void install(mlt_filter f, process_fn ours)
{
    if (!f || !f->process || f->process == rescale_process || f->process == resize_process || f->process == mirror_process)
        return;
    mlt_properties_set_data(MLT_FILTER_PROPERTIES(f), "_cuda_stock_process", (void *) f->process, 0, NULL, NULL);
    f->process = ours;
}

void cuda_rescale_install(mlt_filter stock_rescale)
{
    /* "filter-create-done" fires inside mlt_factory_filter(profile, "rescale"): "method" is still core's filter_scale */
    if (stock_rescale && stock_rescale->process != rescale_process)
        mlt_properties_set_data(MLT_FILTER_PROPERTIES(stock_rescale), CUDA_CORE_SCALER,
                                mlt_properties_get_data(MLT_FILTER_PROPERTIES(stock_rescale), "method", NULL), 0, NULL, NULL);
    install(stock_rescale, rescale_process);
}

void cuda_resize_install(mlt_filter stock_resize)
{
    install(stock_resize, resize_process);
}

void cuda_mirror_install(mlt_filter stock_mirror)
{
    install(stock_mirror, mirror_process);
}
