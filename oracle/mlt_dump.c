/*
 * mlt_dump.c -- TEST INFRASTRUCTURE.  Renders single frames of a melt-style graph through the
 * public MLT C API of whatever libmlt-7 / modules the environment points at (oracle/_ref for the
 * reference; the same plus mlt_b200/lib/mlt/libmltcuda.so for the drop-in) and dumps raw bytes.
 *
 *   mlt_dump <profile> <format> <out-prefix> <frame>[,<frame>...] -- <melt args...>
 *
 * For each frame N writes <out-prefix>.N.img (image bytes) and <out-prefix>.N.alpha (if the frame
 * has an alpha plane) and prints "N width height format image_fnv1a64 alpha_fnv1a64".
 * Pull sequence as in SURVEY App. C: mlt_producer_seek -> mlt_service_get_frame -> frame props
 * consumer.rescale=bilinear, consumer.progressive=1 -> mlt_frame_get_image.
 */
#include <framework/mlt.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static unsigned long long fnv1a(const unsigned char *p, size_t n)
{
    unsigned long long h = 1469598103934665603ULL;
    for (size_t i = 0; i < n; i++) {
        h ^= p[i];
        h *= 1099511628211ULL;
    }
    return h;
}

/* MLT_DUMP_FOREIGN_RESCALER=1: stand-in for swscale / gtkrescale, which this build lacks.  Like them
 * (src/modules/avformat/filter_swscale.c:391-402, src/modules/gdk/filter_rescale.c:140-151) it is a filter that
 * creates core `rescale` by name and then swaps the scaler function ("method") for its own; registered as
 * "gtkrescale", which loader.ini prefers over "rescale".  The scaler is deliberately unlike core's: it samples
 * whole units (a yuv422 pair, an rgb / rgba pixel) at the unit centres and flips one bit of each unit's first
 * byte, so a picture that went through core's scaler -- or the device copy of it -- instead cannot match. */
static int toy_scale(mlt_frame frame, uint8_t **image, mlt_image_format *format, int iw, int ih, int ow, int oh)
{
    const int unit = *format == mlt_image_rgb ? 3 : 4, ppu = *format == mlt_image_yuv422 ? 2 : 1;
    if (*format != mlt_image_yuv422 && *format != mlt_image_rgb && *format != mlt_image_rgba)
        return 1;
    const int iu = iw / ppu, ou = ow / ppu;
    const int size = mlt_image_format_size(*format, ow, oh + 1, NULL);
    uint8_t *out = mlt_pool_alloc(size);
    memset(out, 0, size);
    for (int y = 0; y < oh; y++) {
        const uint8_t *in_row = *image + (size_t) ((2 * y + 1) * ih / (2 * oh)) * iw * unit / ppu;
        uint8_t *o = out + (size_t) y * ow * unit / ppu;
        for (int x = 0; x < ou; x++, o += unit) {
            memcpy(o, in_row + (size_t) ((2 * x + 1) * iu / (2 * ou)) * unit, unit);
            o[0] ^= 0x20;
        }
    }
    mlt_frame_set_image(frame, out, size, mlt_pool_release);
    *image = out;
    return 0;
}

static void *toy_rescaler_init(mlt_profile profile, mlt_service_type type, const char *id, const void *arg)
{
    mlt_filter filter = mlt_factory_filter(profile, "rescale", arg);
    if (filter)
        mlt_properties_set_data(MLT_FILTER_PROPERTIES(filter), "method", toy_scale, 0, NULL, NULL);
    return filter;
}

int main(int argc, char **argv)
{
    if (argc < 7 || strcmp(argv[5], "--")) {
        fprintf(stderr, "usage: %s profile format prefix frames -- melt-args\n", argv[0]);
        return 2;
    }
    const char *rescale = getenv("MLT_DUMP_RESCALE") ? getenv("MLT_DUMP_RESCALE") : "bilinear";
    mlt_factory_init(NULL);
    if (getenv("MLT_DUMP_FOREIGN_RESCALER"))
        mlt_repository_register(mlt_factory_repository(), mlt_service_filter_type, "gtkrescale", toy_rescaler_init);
    mlt_profile profile = mlt_profile_init(argv[1]);
    if (!profile) {
        fprintf(stderr, "no profile %s\n", argv[1]);
        return 2;
    }
    profile->is_explicit = 1;
    mlt_image_format want = mlt_image_format_id(argv[2]);
    mlt_producer melt = mlt_factory_producer(profile, "melt", &argv[6]);
    if (!melt) {
        fprintf(stderr, "melt producer failed\n");
        return 2;
    }
    char *frames = strdup(argv[4]);
    for (char *tok = strtok(frames, ","); tok; tok = strtok(NULL, ",")) {
        int n = atoi(tok);
        mlt_producer_seek(melt, n);
        mlt_frame frame = NULL;
        if (mlt_service_get_frame(MLT_PRODUCER_SERVICE(melt), &frame, 0) || !frame) {
            fprintf(stderr, "get_frame %d failed\n", n);
            return 3;
        }
        mlt_properties fp = MLT_FRAME_PROPERTIES(frame);
        mlt_properties_set(fp, "consumer.rescale", rescale);
        mlt_properties_set_int(fp, "consumer.progressive", 1);
        uint8_t *image = NULL;
        mlt_image_format fmt = want;
        int w = profile->width, h = profile->height;
        int err = mlt_frame_get_image(frame, &image, &fmt, &w, &h, 0);
        if (err || !image) {
            fprintf(stderr, "get_image %d failed (%d)\n", n, err);
            return 4;
        }
        int size = mlt_image_format_size(fmt, w, h, NULL);
        uint8_t *alpha = mlt_frame_get_alpha(frame);
        char path[4096];
        snprintf(path, sizeof(path), "%s.%d.img", argv[3], n);
        FILE *f = fopen(path, "wb");
        fwrite(image, 1, size, f);
        fclose(f);
        if (alpha) {
            snprintf(path, sizeof(path), "%s.%d.alpha", argv[3], n);
            f = fopen(path, "wb");
            fwrite(alpha, 1, (size_t) w * h, f);
            fclose(f);
        }
        printf("%d %d %d %s %016llx %016llx\n", n, w, h, mlt_image_format_name(fmt), fnv1a(image, size),
               alpha ? fnv1a(alpha, (size_t) w * h) : 0ULL);
        mlt_frame_close(frame);
    }
    mlt_producer_close(melt);
    mlt_profile_close(profile);
    mlt_factory_close();
    return 0;
}
