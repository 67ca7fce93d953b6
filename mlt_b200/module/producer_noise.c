/*
 * producer_noise.c (cuda module) -- the `noise` producer with its picture generated on the GPU.
 *
 * Replaces src/modules/core/producer_noise.c:57-205 of the reference: same
 * service name, same frame properties, same pixels.  The picture comes from mltcuda_noise()
 * (K6 k_noise: each thread jumps into the reference's two 16-bit multiply-with-carry sequences
 * and reproduces them exactly, buffer filled backwards, :108-123), so a CUDA transition that pulls
 * the frame gets a device-resident image without any host->device copy, and a host consumer gets
 * a download of 2 B/px instead of ~25 ms of scalar RNG per UHD frame.  Audio (white noise from the
 * same generator, :128-170) is a few thousand samples and stays on the host.
 */
#include "cuda_module.h"

#include <stdlib.h>

/* the generator of producer_noise.c:31-51, for the audio samples */
typedef struct
{
    uint32_t x, y;
} noise_rng;

static void noise_seed(noise_rng *r, int position)
{
    r->x = 521288629u + (uint32_t) position - ((uint32_t) position << 16);
    r->y = 362436069u - (uint32_t) position + ((uint32_t) position << 16);
}

static uint32_t noise_next(noise_rng *r)
{
    r->x = 18000u * (r->x & 65535u) + (r->x >> 16);
    r->y = 30903u * (r->y & 65535u) + (r->y >> 16);
    return (r->x << 16) + (r->y & 65535u);
}

static int noise_get_image_impl(mlt_frame frame, uint8_t **buffer, mlt_image_format *format, int *width, int *height,
                           int writable)
{
    mlt_profile profile = mlt_service_profile(MLT_PRODUCER_SERVICE(mlt_frame_get_original_producer(frame)));
    if (*width <= 0)
        *width = profile->width;
    if (*height <= 0)
        *height = profile->height;
    const int resident = cuda_request_resident(frame, *format);

    mltcuda_ctx *ctx = cuda_thread_ctx();
    if (!ctx)
        return 1; /* logged by cuda_thread_ctx: no device, no CPU fallback */
    cuda_image ci = cuda_image_new(ctx, mlt_image_yuv422, *width, *height, 0);
    if (!ci)
        return 1;
    if (mltcuda_noise(ctx, ci->data, *width, *height, mlt_frame_get_position(frame), 1, 0)) {
        mlt_log_error(NULL, "[cuda] noise kernel failed: %s\n", mltcuda_last_error());
        cuda_image_close(ci);
        return 1;
    }
    if (cuda_frame_set_image(frame, ci)) {
        cuda_image_close(ci);
        return 1;
    }
    *buffer = (uint8_t *) ci, *format = mlt_image_private;
    if (resident)
        return 0;
    /* a host filter or the consumer is asking: hand over an ordinary yuv422 image */
    return cuda_download_to_frame(frame, ci, buffer, format);
}

static int noise_get_image(mlt_frame frame, uint8_t **buffer, mlt_image_format *format, int *width, int *height, int writable)
{
    const uint64_t st0 = CUDA_STAT_T0();
    int rc = noise_get_image_impl(frame, buffer, format, width, height, writable);
    CUDA_STAT_END(CS_NOISE, st0);
    return rc;
}

static int noise_get_audio(mlt_frame frame, int16_t **buffer, mlt_audio_format *format, int *frequency, int *channels,
                           int *samples)
{
    *samples = *samples <= 0 ? 1920 : *samples;
    *channels = *channels <= 0 ? 2 : *channels;
    *frequency = *frequency <= 0 ? 48000 : *frequency;
    *format = mlt_audio_s16;
    const int count = *samples * *channels;
    const int size = count * (int) sizeof(int16_t);
    *buffer = mlt_pool_alloc(size);
    if (*buffer) {
        noise_rng rng;
        noise_seed(&rng, mlt_frame_get_position(frame));
        for (int i = count - 1; i >= 0; i--) /* filled from the end, like the picture */
            (*buffer)[i] = (int16_t) noise_next(&rng);
    }
    mlt_frame_set_audio(frame, *buffer, *format, size, mlt_pool_release);
    mlt_properties fp = MLT_FRAME_PROPERTIES(frame);
    mlt_properties_set(fp, "channel_layout", mlt_properties_get(fp, "consumer.channel_layout"));
    return 0;
}

static int noise_get_frame(mlt_producer producer, mlt_frame_ptr frame, int index)
{
    *frame = mlt_frame_init(MLT_PRODUCER_SERVICE(producer));
    if (*frame) {
        mlt_properties fp = MLT_FRAME_PROPERTIES(*frame);
        mlt_properties_set_double(fp, "aspect_ratio", mlt_profile_sar(mlt_service_profile(MLT_PRODUCER_SERVICE(producer))));
        mlt_properties_set_int(fp, "progressive", 1);
        mlt_properties_set_int(fp, "format", mlt_image_yuv422);
        mlt_frame_set_position(*frame, mlt_producer_position(producer));
        mlt_frame_push_get_image(*frame, noise_get_image);
        mlt_frame_push_audio(*frame, noise_get_audio);
    }
    mlt_producer_prepare_next(producer);
    return 0;
}

static void noise_close(mlt_producer producer)
{
    producer->close = NULL;
    mlt_producer_close(producer);
    free(producer);
}

mlt_producer cuda_producer_noise_init(mlt_profile profile, mlt_service_type type, const char *id, char *arg)
{
    mlt_producer producer = mlt_producer_new(profile);
    if (producer) {
        producer->get_frame = noise_get_frame;
        producer->close = (mlt_destructor) noise_close;
        mlt_properties_set_int(MLT_PRODUCER_PROPERTIES(producer), "_cuda_producer", 1);
    }
    return producer;
}
