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

int main(int argc, char **argv)
{
    if (argc < 7 || strcmp(argv[5], "--")) {
        fprintf(stderr, "usage: %s profile format prefix frames -- melt-args\n", argv[0]);
        return 2;
    }
    const char *rescale = getenv("MLT_DUMP_RESCALE") ? getenv("MLT_DUMP_RESCALE") : "bilinear";
    mlt_factory_init(NULL);
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
