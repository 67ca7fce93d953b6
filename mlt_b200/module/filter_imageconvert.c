/*
 * filter_imageconvert.c (cuda module) -- the `imageconvert` filter on the GPU.
 *
 * Drop-in for src/modules/core/filter_imageconvert.c of the reference: the filter's process()
 * registers a converter on every frame (:498-502).  Here the converter is cuda_convert_image
 * (cuda_image.c): the same conversion matrix (:423-437) evaluated by the K1 kernels, with the
 * reference's alpha-plane rules (:462-480), on device-resident or host images.  Like the reference
 * it is APPENDED (mlt_frame_append_convert_image, :500): in a build where avcolor_space precedes
 * imageconvert (core/loader.ini:10) that converter keeps the first try at host conversions.  The
 * CUDA transitions additionally prepend the residency converter (cuda_convert_image), which only
 * ever touches device-resident images.
 */
#include "cuda_module.h"

static mlt_frame imageconvert_process(mlt_filter filter, mlt_frame frame)
{
    cuda_imageconvert_attach(frame);
    return frame;
}

mlt_filter cuda_filter_imageconvert_init(mlt_profile profile, mlt_service_type type, const char *id, char *arg)
{
    mlt_filter filter = mlt_filter_new();
    if (filter)
        filter->process = imageconvert_process;
    return filter;
}
