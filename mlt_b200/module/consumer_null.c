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
 *   - on "consumer-frame-render" (fired by the worker right before it pulls, :1155) it works out
 *     whether its own request is the only host-format request the rendered picture can meet
 *     (consumer_is_sole_requester below) and, if so, puts the one-shot "_cuda_handoff" mark on the
 *     frame the module's converter will see.  mlt_image_private cannot be requested through
 *     "mlt_image_format" (it sorts after mlt_image_invalid, mlt_types.h:52-53, so
 *     mlt_image_format_id never returns it): the workers keep asking for the host format, and the
 *     converter, where the download would happen, honours the mark instead -- it runs the deferred
 *     op list (one fused launch, in the worker thread), records an event on the image and returns
 *     the frame still in format mlt_image_private;
 *   - on "consumer-frame-show" (the consumer thread, consumer_null.c:153) the sink waits for that
 *     event: the frame's pixels are final in HBM, nothing was copied.
 * Any other consumer can follow the same three steps.  A frame that reaches it with a host image
 * (no CUDA service on the graph, or a CPU service above the last CUDA one: then the mark is not
 * set and the converter downloads as for any host consumer) is simply passed through.
 */
#include "cuda_module.h"

/* Will the worker's mlt_frame_get_image(frame, <host format>) be the FIRST host-format request to reach the
 * picture the CUDA services render?  Only then may the converter answer it with the device handle: a stock CPU
 * service asking in between (a filter attached to the tractor, a filter planted or a CPU transition placed
 * after the last CUDA transition) expects pixels and would read the ~64-byte handle as width*height of them.
 * What can be proved from the stacks as the graph walk left them:
 *   - the consumer pulls from a tractor and the output frame's image stack holds exactly the two entries
 *     mlt_tractor pushes (the track frame and its producer_get_image, mlt_tractor.c:596-597): nothing
 *     attached to the tractor sits above them;
 *   - on that track frame the top entry -- the callback the tractor's request runs first -- is a get_image of
 *     one of this module's transitions, which pull their own inputs with "_cuda_requested" raised.
 * The track frame is returned (it is where the converter runs); NULL = hand nothing off. */
static mlt_frame consumer_is_sole_requester(mlt_consumer self, mlt_frame frame)
{
    mlt_service producer = mlt_service_producer(MLT_CONSUMER_SERVICE(self));
    if (!producer || mlt_service_identify(producer) != mlt_service_tractor_type)
        return NULL;
    mlt_deque stack = MLT_FRAME_IMAGE_STACK(frame);
    if (mlt_deque_count(stack) != 2)
        return NULL;
    mlt_frame track = mlt_deque_peek(stack, 0);
    stack = track ? MLT_FRAME_IMAGE_STACK(track) : NULL;
    if (!stack || mlt_deque_count(stack) < 1 || !cuda_is_module_get_image(mlt_deque_peek_back(stack)))
        return NULL;
    return track;
}

static void on_frame_render(mlt_properties owner, mlt_consumer self, mlt_event_data ev)
{
    mlt_frame frame = mlt_event_data_to_frame(ev);
    mlt_frame track = frame ? consumer_is_sole_requester(self, frame) : NULL;
    if (!track)
        return;
    mlt_properties_set_int(MLT_FRAME_PROPERTIES(track), CUDA_HANDOFF, 1);
    mlt_properties_set_int(MLT_FRAME_PROPERTIES(frame), CUDA_CONSUMER, 1); /* for on_frame_show */
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
    if (mlt_properties_get_int(p, CUDA_CONSUMER) > 0 && mlt_properties_get_int(p, "format") == mlt_image_private) {
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
