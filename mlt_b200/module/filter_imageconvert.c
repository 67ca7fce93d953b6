/*
 * filter_imageconvert.c (cuda module) -- the `imageconvert` filter on the GPU.
 *
 * Drop-in for src/modules/core/filter_imageconvert.c of the reference: the filter's process()
 * registers a converter on every frame (:498-502).  Here the converter is cuda_convert_image
 * (cuda_image.c): the same conversion matrix (:423-437) evaluated by the K1 kernels, with the
 * reference's alpha-plane rules (:462-480), on device-resident or host images.  It is prepended
 * rather than appended (mlt_frame_prepend_convert_image, src/framework/mlt_frame.c:915-935) so it
 * also wins in builds where avcolor_space precedes imageconvert in core/loader.ini:10.
 */
#include "cuda_module.h"

static mlt_frame imageconvert_process(mlt_filter filter, mlt_frame frame)
{
    mlt_frame_prepend_convert_image(frame, cuda_convert_image);
    return frame;
}

mlt_filter cuda_filter_imageconvert_init(mlt_profile profile, mlt_service_type type, const char *id, char *arg)
{
    mlt_filter filter = mlt_filter_new();
    if (filter)
        filter->process = imageconvert_process;
    return filter;
}
