/*
 * consumer_null.c (cuda module) -- `cuda.null`: the null consumer that takes device-resident frames
 * (SURVEY 8(f) f4).
 *
 * The reference's null consumer (src/modules/core/consumer_null.c:130-167) pulls rendered frames
 * from mlt_consumer_rt_frame and drops them; the render itself is the mlt_frame_get_image call of
 * the consumer's worker threads (src/framework/mlt_consumer.c:1158), asked for the consumer's
 * "mlt_image_format".  Asked for a host format, every frame costs a device->host copy (25 MB at
 * 2160p with its alpha plane) that a sink living on the GPU -- an encoder, a device-side analyser --
 * has no use for.
 *
 * This service is the stock `null` consumer with the hand-off contract such a sink follows:
 *   - on "consumer-frame-render" (fired by the worker right before it pulls, :1155) the frame is
 *     marked "consumer._cuda_consumer" (mlt_tractor copies "consumer." properties down to the track
 *     frames, src/framework/mlt_tractor.c:372).  mlt_image_private cannot be requested through
 *     "mlt_image_format" (it sorts after mlt_image_invalid, mlt_types.h:52-53, so
 *     mlt_image_format_id never returns it): the workers keep asking for the host format, and the
 *     module's converter, where the download would happen, honours the mark instead -- it runs the
 *     deferred op list (one fused launch, in the worker thread), records an event on the image and
 *     returns the frame still in format mlt_image_private;
 *   - on "consumer-frame-show" (the consumer thread, consumer_null.c:153) the sink waits for that
 *     event: the frame's pixels are final in HBM, nothing was copied.
 * Any other consumer can follow the same three steps.  A frame that reaches it with a host image
 * (no CUDA service on the graph) or through a path that dropped the handle is simply passed through.
 */
#include "cuda_module.h"

static void on_frame_render(mlt_properties owner, mlt_consumer self, mlt_event_data ev)
{
    mlt_frame frame = mlt_event_data_to_frame(ev);
    if (frame) /* "consumer." properties reach every track frame (mlt_tractor.c:372) */
        mlt_properties_set_int(MLT_FRAME_PROPERTIES(frame), "consumer." CUDA_CONSUMER, 1);
}

static void on_frame_show(mlt_properties owner, mlt_consumer self, mlt_event_data ev)
{
    mlt_frame frame = mlt_event_data_to_frame(ev);
    if (!frame)
        return;
    mlt_properties p = MLT_FRAME_PROPERTIES(frame);
    /* the tractor's output frame carries the track frame's handle but not its tag property
     * (mlt_tractor.c:381-386): recognise the handle by its magic, on frames this consumer marked */
    cuda_image ci = NULL;
    if (mlt_properties_get_int(p, "consumer." CUDA_CONSUMER) > 0 && mlt_properties_get_int(p, "format") == mlt_image_private) {
        ci = mlt_properties_get_data(p, "image", NULL);
        if (ci && ci->magic != CUDA_IMAGE_MAGIC)
            ci = NULL;
    }
    if (ci && cuda_image_wait_ready(ci))
        mlt_log_error(MLT_CONSUMER_SERVICE(self), "[cuda] frame %d did not complete: %s\n", (int) mlt_frame_get_position(frame),
                      mltcuda_last_error());
}

mlt_consumer cuda_consumer_null_init(mlt_profile profile, mlt_service_type type, const char *id, char *arg)
{
    mlt_consumer consumer = mlt_factory_consumer(profile, "null", arg);
    if (!consumer)
        return NULL;
    mlt_properties props = MLT_CONSUMER_PROPERTIES(consumer);
    mlt_events_listen(props, consumer, "consumer-frame-render", (mlt_listener) on_frame_render);
    mlt_events_listen(props, consumer, "consumer-frame-show", (mlt_listener) on_frame_show);
    return consumer;
}
