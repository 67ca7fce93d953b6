/*
 * factory.c (cuda module) -- mlt_register for libmltcuda.so.
 *
 * Registers the four services of the compositing hot path under the reference's names
 * (`composite`, `luma`, `affine` transitions: src/modules/core/factory.c:234-236,
 * src/modules/plus/factory.c:190; `imageconvert` filter: core/factory.c:231) and under `cuda.*`
 * aliases so both implementations can be instantiated side by side.
 *
 * Deterministic override: the repository loads modules in readdir order and a duplicate name
 * simply overwrites (src/framework/mlt_repository.c:79,263-276), so registration alone cannot
 * guarantee which `composite` wins.  The module therefore also listens to the factory's
 * "transition-create-request" / "filter-create-request" events (src/framework/mlt_factory.c:
 * 370-388,425-443): when the requested name is one of ours it creates the service itself and the
 * factory skips the repository.  MLT_CUDA_OVERRIDE=0 turns this off (aliases only).
 */
#include "cuda_module.h"

#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static const char *base_name(const char *id)
{
    return id && !strncmp(id, "cuda.", 5) ? id + 5 : id;
}

static void *create_transition(mlt_profile profile, mlt_service_type type, const char *id, const void *arg)
{
    const char *n = base_name(id);
    if (!strcmp(n, "composite"))
        return cuda_transition_composite_init(profile, type, id, (char *) arg);
    if (!strcmp(n, "luma"))
        return cuda_transition_luma_init(profile, type, id, (char *) arg);
    if (!strcmp(n, "affine"))
        return cuda_transition_affine_init(profile, type, id, (char *) arg);
    return NULL;
}

static void *create_filter(mlt_profile profile, mlt_service_type type, const char *id, const void *arg)
{
    if (!strcmp(base_name(id), "imageconvert"))
        return cuda_filter_imageconvert_init(profile, type, id, (char *) arg);
    return NULL;
}

static void *create_producer(mlt_profile profile, mlt_service_type type, const char *id, const void *arg)
{
    const char *n = base_name(id);
    if (!strcmp(n, "noise"))
        return cuda_producer_noise_init(profile, type, id, (char *) arg);
    if (!strcmp(n, "colour") || !strcmp(n, "color"))
        return cuda_producer_colour_init(profile, type, id, (char *) arg);
    return NULL;
}

static void *create_consumer(mlt_profile profile, mlt_service_type type, const char *id, const void *arg)
{
    return cuda_consumer_null_init(profile, type, id, (char *) arg);
}

static int override_enabled(void)
{
    const char *e = getenv("MLT_CUDA_OVERRIDE");
    return !(e && (!strcmp(e, "0") || !strcmp(e, "off") || !strcmp(e, "no")));
}

static void on_transition_request(mlt_properties owner, void *unused, mlt_event_data ev)
{
    mlt_factory_event_data *d = mlt_event_data_to_object(ev);
    if (!d || !d->name || !d->service || *(void **) d->service || !override_enabled())
        return;
    if (!strcmp(d->name, "composite") || !strcmp(d->name, "luma") || !strcmp(d->name, "affine"))
        *(void **) d->service = create_transition(NULL, mlt_service_transition_type, d->name, d->input);
}

static void on_filter_request(mlt_properties owner, void *unused, mlt_event_data ev)
{
    mlt_factory_event_data *d = mlt_event_data_to_object(ev);
    if (!d || !d->name || !d->service || *(void **) d->service || !override_enabled())
        return;
    if (!strcmp(d->name, "imageconvert"))
        *(void **) d->service = create_filter(NULL, mlt_service_filter_type, d->name, d->input);
}

/* producers: MLT_CUDA_PRODUCERS=0 keeps the stock CPU producers while the
 * transitions stay overridden */
static int producers_enabled(void)
{
    const char *e = getenv("MLT_CUDA_PRODUCERS");
    return override_enabled() && !(e && (!strcmp(e, "0") || !strcmp(e, "off") || !strcmp(e, "no")));
}

/* "producer-create-request" carries no profile and the factory does not stamp one on a producer a
 * listener supplies (mlt_factory.c:337-358), so the override hooks "producer-create-done" instead:
 * the repository has just built the `noise` producer -- ours if this module registered the name
 * last, the stock one otherwise (load order is readdir order, mlt_repository.c:79) -- and a stock
 * one is swapped for ours built on the same profile before the factory stamps the common
 * properties on it. */
static void on_producer_done(mlt_properties owner, void *unused, mlt_event_data ev)
{
    mlt_factory_event_data *d = mlt_event_data_to_object(ev);
    if (!d || !d->name || !d->service || !*(void **) d->service || !producers_enabled())
        return;
    if (cuda_creating_stock_producer || (strcmp(d->name, "noise") && strcmp(d->name, "colour") && strcmp(d->name, "color")))
        return;
    mlt_producer stock = *(mlt_producer *) d->service;
    if (mlt_properties_get_int(MLT_PRODUCER_PROPERTIES(stock), "_cuda_producer"))
        return;
    mlt_producer ours = create_producer(mlt_service_profile(MLT_PRODUCER_SERVICE(stock)), mlt_service_producer_type, d->name,
                                        d->input);
    if (ours) {
        mlt_producer_close(stock);
        *(mlt_producer *) d->service = ours;
    }
}

/* MLT_CUDA_MIRROR=0 leaves core `mirror` on the host (a device producer's frame is then downloaded for it) */
static int mirror_enabled(void)
{
    const char *e = getenv("MLT_CUDA_MIRROR");
    return !(e && (!strcmp(e, "0") || !strcmp(e, "off") || !strcmp(e, "no")));
}

/* the loader's stock `rescale` / `resize` normalisers get their device paths (filter_rescale.c) */
static void on_filter_done(mlt_properties owner, void *unused, mlt_event_data ev)
{
    mlt_factory_event_data *d = mlt_event_data_to_object(ev);
    if (!d || !d->name || !d->service || !*(void **) d->service || !producers_enabled())
        return;
    if (!strcmp(d->name, "rescale"))
        cuda_rescale_install(*(mlt_filter *) d->service);
    if (!strcmp(d->name, "resize"))
        cuda_resize_install(*(mlt_filter *) d->service);
    if (!strcmp(d->name, "mirror") && mirror_enabled())
        cuda_mirror_install(*(mlt_filter *) d->service);
}

static mlt_properties metadata(mlt_service_type type, const char *id, void *data)
{
    char file[PATH_MAX];
    snprintf(file, sizeof(file), "%s/cuda/%s", mlt_environment("MLT_DATA"), (char *) data);
    return mlt_properties_parse_yaml(file);
}

MLT_REPOSITORY
{
    static const char *transitions[] = {"composite", "luma", "affine"};
    static char yml[4][64];
    for (int i = 0; i < 3; i++) {
        char alias[32];
        snprintf(alias, sizeof(alias), "cuda.%s", transitions[i]);
        snprintf(yml[i], sizeof(yml[i]), "transition_%s.yml", transitions[i]);
        MLT_REGISTER(mlt_service_transition_type, alias, create_transition);
        MLT_REGISTER_METADATA(mlt_service_transition_type, alias, metadata, yml[i]);
        if (override_enabled()) {
            MLT_REGISTER(mlt_service_transition_type, transitions[i], create_transition);
            MLT_REGISTER_METADATA(mlt_service_transition_type, transitions[i], metadata, yml[i]);
        }
    }
    snprintf(yml[3], sizeof(yml[3]), "filter_imageconvert.yml");
    MLT_REGISTER(mlt_service_filter_type, "cuda.imageconvert", create_filter);
    MLT_REGISTER_METADATA(mlt_service_filter_type, "cuda.imageconvert", metadata, yml[3]);
    if (override_enabled()) {
        MLT_REGISTER(mlt_service_filter_type, "imageconvert", create_filter);
        MLT_REGISTER_METADATA(mlt_service_filter_type, "imageconvert", metadata, yml[3]);
    }
    /* link flavour, as the reference registers it (core/factory.c:248): mlt_link_filter_init wraps the
     * filter of the same name, which the lines above resolve to this module */
    MLT_REGISTER(mlt_service_link_type, "cuda.imageconvert", mlt_link_filter_init);
    if (override_enabled())
        MLT_REGISTER(mlt_service_link_type, "imageconvert", mlt_link_filter_init);
    MLT_REGISTER(mlt_service_consumer_type, "cuda.null", create_consumer);
    MLT_REGISTER_METADATA(mlt_service_consumer_type, "cuda.null", metadata, "consumer_null.yml");
    MLT_REGISTER(mlt_service_producer_type, "cuda.colour", create_producer);
    MLT_REGISTER_METADATA(mlt_service_producer_type, "cuda.colour", metadata, "producer_colour.yml");
    MLT_REGISTER(mlt_service_producer_type, "cuda.noise", create_producer);
    MLT_REGISTER_METADATA(mlt_service_producer_type, "cuda.noise", metadata, "producer_noise.yml");
    if (producers_enabled()) {
        MLT_REGISTER(mlt_service_producer_type, "noise", create_producer);
        MLT_REGISTER_METADATA(mlt_service_producer_type, "noise", metadata, "producer_noise.yml");
        /* `colour` / `color` are NOT re-registered: the wrapper instantiates the stock producer by
         * name for everything off the fast path, and is swapped in by on_producer_done */
    }
    mlt_properties events = mlt_factory_event_object();
    if (events) {
        mlt_events_listen(events, NULL, "transition-create-request", (mlt_listener) on_transition_request);
        mlt_events_listen(events, NULL, "filter-create-request", (mlt_listener) on_filter_request);
        mlt_events_listen(events, NULL, "producer-create-done", (mlt_listener) on_producer_done);
        mlt_events_listen(events, NULL, "filter-create-done", (mlt_listener) on_filter_done);
    }
}
