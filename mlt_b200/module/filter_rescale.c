/* filter_rescale.c -- residency shim for the loader's stock `rescale` normaliser.
 *
 * The loader attaches rescale + resize (+ crop, fieldorder, deinterlace) to every producer
 * (src/modules/core/loader.ini, producer_loader.c:232-248).  With its built-in scaler the stock
 * rescale filter overwrites the requested format with yuv422 before pulling its input
 * (src/modules/core/filter_rescale.c:221-223), so a device producer below it never sees the
 * mlt_image_private request of a CUDA transition: the picture was downloaded, handed up as a host
 * image and uploaded again -- two 16.6 MB PCIe trips per 2160p frame, the melt-level bottleneck.
 *
 * The stock filter is left in place (same property bookkeeping, same scaling of host images); this
 * shim sits right above it on the frame's get_image stack.  When a CUDA service is pulling and
 * nothing above or below will touch pixels (no crop, no padding by resize), it marks the frame
 * CUDA_PASSTHROUGH for the duration of the nested call: device producers then answer the forced
 * yuv422 request with a resident yuv422 image, and cuda_convert_image leaves it resident on the
 * way back up.  Host images take the stock path unchanged.  Sizes: device producers render at the
 * requested size, so the stock filter finds nothing to scale. */
#include "cuda_module.h"

#include <string.h>

typedef mlt_frame (*process_fn)(mlt_filter, mlt_frame);

static int shim_get_image(mlt_frame frame, uint8_t **image, mlt_image_format *format, int *width, int *height, int writable)
{
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    int pass = *format == mlt_image_private && mlt_properties_get_int(p, CUDA_REQUESTED) > 0;
    /* crop would cut host pixels (filter_crop.c:74-77) */
    if (pass
        && (mlt_properties_get_int(p, "crop.left") || mlt_properties_get_int(p, "crop.right") || mlt_properties_get_int(p, "crop.top")
            || mlt_properties_get_int(p, "crop.bottom")))
        pass = 0;
    /* resize pads whatever comes back smaller than it was asked for (filter_resize.c:143-193) */
    const int rw = mlt_properties_get_int(p, "resize_width"), rh = mlt_properties_get_int(p, "resize_height");
    if (pass && (rw || rh) && (rw != *width || rh != *height))
        pass = 0;
    const int saved = mlt_properties_get_int(p, CUDA_PASSTHROUGH);
    mlt_properties_set_int(p, CUDA_PASSTHROUGH, pass);
    int err = mlt_frame_get_image(frame, image, format, width, height, writable); /* the stock rescale */
    mlt_properties_set_int(p, CUDA_PASSTHROUGH, saved);
    return err;
}

static mlt_frame shim_process(mlt_filter filter, mlt_frame frame)
{
    process_fn stock = (process_fn) mlt_properties_get_data(MLT_FILTER_PROPERTIES(filter), "_cuda_stock_process", NULL);
    if (stock)
        frame = stock(filter, frame);
    mlt_frame_push_get_image(frame, shim_get_image);
    return frame;
}

void cuda_rescale_install(mlt_filter f)
{
    if (!f || !f->process || f->process == shim_process)
        return;
    mlt_properties_set_data(MLT_FILTER_PROPERTIES(f), "_cuda_stock_process", (void *) f->process, 0, NULL, NULL);
    f->process = shim_process;
}
