/*
 * mlt_oracle.c -- TEST INFRASTRUCTURE.  CPU restatement (plain C, strict IEEE: build with
 * -O2 -ffp-contract=off, no -ffast-math) of the reference's compositing hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (mlt_b200/) never links or calls it.
 *
 * Parity pin: checked against the UNMODIFIED reference compiled into oracle/_ref (Makefile.ref)
 * by tests/test_oracle_pinning.py and tests/test_cfg_vs_reference.py (live, when oracle/_ref
 * exists) and against the committed fixtures tests/golden/ (npz files, frames.json) that
 * tests/golden/make_golden.py produced by calling that build (composite_line_yuv,
 * interp*_b32/_b64 via dlopen; imageconvert / luma / affine / composite frames via the public MLT
 * API, oracle/mlt_dump.c).  The float luma paths (translucent wipes and dissolves, rgba64) are
 * pinned where they run: tests/test_module.py renders them with the stock services and with the
 * CUDA services inside the reference's own melt (+-1 LSB).  The reference's own test-suite holds no vectors for this path
 * (SURVEY.md 4, 8c).  rgba64 colour conversion is defined by this project ("parity unpinned":
 * libswscale, not in the reference tree).
 *
 * Every function cites the reference file:line (relative to the reference root) it follows.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ---------------------------------------------------------------------------------------- */
/* composite: src/modules/core/transition_composite.c:112-275 and
 * src/modules/core/composite_line_yuv_sse2_simple.c:29-244                                  */

/* transition_composite.c:112-123 (and the identical copy transition_luma.c:297-308) */
static inline int32_t orc_smoothstep(int32_t edge1, int32_t edge2, uint32_t a)
{
    if (a < (uint32_t) edge1)
        return 0;
    if (a >= (uint32_t) edge2)
        return 0x10000;
    a = ((a - (uint32_t) edge1) << 16) / (uint32_t) (edge2 - edge1);
    return (int32_t) ((((a * a) >> 16) * ((3u << 16) - (2 * a))) >> 16);
}

/* one 16-bit lane of DST_PIX_CALC / DST_A_CALC: ((v>>8)+v+128)>>8 in uint16 arithmetic
 * (composite_line_yuv_sse2_simple.c:56-101) */
static inline uint16_t orc_div255(uint16_t v)
{
    return (uint16_t) ((uint16_t) ((uint16_t) (v >> 8) + v + 128) >> 8);
}

/* op: 0 over, 1 or, 2 and, 3 xor (transition_composite.c:143-275).
 * scalar_only: 1 = a build without USE_SSE/ARCH_X86_64 (:156-167 compiled out). */
ORC_API void orc_composite_line_yuv(uint8_t *dest, const uint8_t *src, int width,
                                    const uint8_t *alpha_b, uint8_t *alpha_a, int weight,
                                    const uint16_t *luma, int soft, uint32_t step, int op,
                                    int scalar_only)
{
    int j = 0;
    if (op == 0 && !scalar_only && !luma && width > 7) {
        /* composite_line_yuv_sse2_simple.c:198-244 */
        int w8 = weight >> 8;
        int n = width - width % 8;
        int cond = (alpha_a ? 4 : 0) | (alpha_b ? 2 : 0) | (w8 != 256 ? 1 : 0);
        if (cond == 0 || cond == 4) {
            memcpy(dest, src, 2 * (size_t) width);
            if (cond == 4)
                memset(alpha_a, 0xFF, (size_t) width);
        } else {
            for (int p = 0; p < n; p++) {
                uint16_t A;
                if (alpha_b)
                    A = (cond & 1) ? (uint16_t) ((uint16_t) (alpha_b[p] * w8) >> 8) : alpha_b[p];
                else
                    A = (uint16_t) w8;
                if (alpha_a) {
                    uint16_t da = alpha_a[p];
                    uint16_t t = (uint16_t) ((uint16_t) (255 - da) * A);
                    uint16_t r = (uint16_t) (da + orc_div255(t));
                    alpha_a[p] = r > 255 ? 255 : (uint8_t) r; /* packuswb */
                }
                for (int c = 0; c < 2; c++) {
                    uint16_t d = dest[2 * p + c], s = src[2 * p + c];
                    uint16_t v = (uint16_t) ((uint16_t) (d * 255) + (uint16_t) ((uint16_t) (s - d) * A));
                    uint16_t r = orc_div255(v);
                    /* packuswb: signed 16 -> unsigned 8 saturation */
                    int16_t sr = (int16_t) r;
                    dest[2 * p + c] = sr < 0 ? 0 : (sr > 255 ? 255 : (uint8_t) sr);
                }
            }
        }
        j = n;
    }
    for (; j < width; j++) {
        int ab = alpha_b ? alpha_b[j] : 255;
        int a;
        switch (op) {
        case 1: a = ab | (alpha_a ? alpha_a[j] : 255); break;
        case 2: a = ab & (alpha_a ? alpha_a[j] : 255); break;
        case 3: a = ab ^ (alpha_a ? alpha_a[j] : 255); break;
        default: a = ab; break;
        }
        int base = luma ? orc_smoothstep(luma[j], luma[j] + soft, step) : weight;
        int mix = (base * (a + 1)) >> 8;
        dest[2 * j] = (uint8_t) ((src[2 * j] * mix + dest[2 * j] * ((1 << 16) - mix)) >> 16);
        dest[2 * j + 1] = (uint8_t) ((src[2 * j + 1] * mix + dest[2 * j + 1] * ((1 << 16) - mix)) >> 16);
        if (alpha_a)
            alpha_a[j] = op == 0 ? (uint8_t) ((mix >> 8) | alpha_a[j]) : (uint8_t) (mix >> 8);
    }
}

/* Row loop of composite_yuv (transition_composite.c:455-467). Strides as in mltcuda_composite_desc. */
ORC_API void orc_composite_rows(uint8_t *dest, const uint8_t *src, const uint8_t *alpha_b,
                                uint8_t *alpha_a, const uint16_t *luma, int width, int rows,
                                int stride_dest, int stride_src, int stride_alpha_a,
                                int stride_alpha_b, int weight, int soft, uint32_t step, int op,
                                int scalar_only)
{
    for (int r = 0; r < rows; r++) {
        orc_composite_line_yuv(dest, src, width, alpha_b, alpha_a, weight, luma, soft, step, op,
                               scalar_only);
        dest += stride_dest;
        src += stride_src;
        if (alpha_b)
            alpha_b += stride_alpha_b;
        if (alpha_a)
            alpha_a += stride_alpha_a;
        if (luma)
            luma += stride_alpha_b;
    }
}

/* Whole composite_yuv (transition_composite.c:330-492): geometry clipping + row loop.
 * g = {x,y,w,h,o} item, ints nw,nh,sw,sh,x_src,y_src (struct geometry_s :45-56).
 * Returns number of rows composited. */
ORC_API int orc_composite_yuv(uint8_t *p_dest, int width_dest, int height_dest,
                              const uint8_t *p_src, int width_src, int height_src,
                              const uint8_t *alpha_b, uint8_t *alpha_a, const double *item,
                              const int *gi, int field, const uint16_t *p_luma, double softness,
                              int op, int scalar_only)
{
    double ix = item[0], iy = item[1], iw = item[2], ih = item[3], io = item[4];
    int nw = gi[0], nh = gi[1], sw = gi[2], gx_src = gi[4], gy_src = gi[5];
    int x_src = -gx_src, y_src = -gy_src;
    int uneven_x_src = x_src % 2;
    int step = field > -1 ? 2 : 1;
    int stride_src = sw * 2, stride_dest = width_dest * 2;
    int i_softness = (int) ((1 << 16) * softness);
    int weight = (int) (((1 << 16) * io + 50) / 100);
    uint32_t luma_step = (uint32_t) ((((1 << 16) - 1) * io + 50) / 100 * (1.0 + softness));
    int x = (int) rint(ix * width_dest / nw);
    int y = (int) rint(iy * height_dest / nh);
    int uneven_x = x % 2;

    if (width_src <= 0 || height_src <= 0 || y_src >= height_src || x_src >= width_src)
        return 0;
    if ((x < 0 && -x >= width_src) || (y < 0 && -y >= height_src))
        return 0;
    if (x_src > 0) {
        width_src -= x_src;
        if (width_src > iw)
            width_src = (int) iw;
    }
    if (y_src > 0) {
        height_src -= y_src;
        if (height_src > ih)
            height_src = (int) ih;
    }
    if (x < 0) {
        x_src = -x;
        width_src -= x_src;
        x = 0;
    }
    if (x + width_src > width_dest)
        width_src = width_dest - x;
    if (y < 0) {
        y_src = -y;
        height_src -= y_src;
        y = 0;
    }
    if (y + height_src > height_dest)
        height_src = height_dest - y;

    p_src += x_src * 2 + y_src * stride_src;
    p_dest += x * 2 + y * stride_dest;
    if (alpha_b)
        alpha_b += x_src + y_src * stride_src / 2;
    if (alpha_a)
        alpha_a += x + y * stride_dest / 2;
    if (p_luma)
        p_luma += x_src + y_src * stride_src / 2;

    if (field > -1 && y % 2 == field) {
        if ((field == 1 && y < height_dest - 1) || (field == 0 && y == 0))
            p_dest += stride_dest;
        else
            p_dest -= stride_dest;
    }
    if (field == 1) {
        p_src += stride_src;
        if (alpha_b)
            alpha_b += stride_src / 2;
        if (alpha_a)
            alpha_a += stride_dest / 2;
        height_src--;
    }
    stride_src *= step;
    stride_dest *= step;
    if (uneven_x != uneven_x_src)
        p_src += 2;

    int rows = 0;
    for (int i = 0; i < height_src; i += step)
        rows++;
    orc_composite_rows(p_dest, p_src, alpha_b, alpha_a, p_luma, width_src, rows, stride_dest,
                       stride_src, stride_dest / 2, stride_src / 2, weight, i_softness, luma_step,
                       op, scalar_only);
    return rows;
}

/* scale_luma: transition_composite.c:497-520 */
ORC_API void orc_scale_luma(uint16_t *dst, int dw, int dh, const uint16_t *src, int sw, int sh,
                            int invert_xor)
{
    int x_step = (sw << 16) / dw, y_step = (sh << 16) / dh, y = 0;
    for (int i = 0; i < dh; i++) {
        const uint16_t *row = src + (y >> 16) * sw;
        int x = 0;
        for (int j = 0; j < dw; j++) {
            *dst++ = row[x >> 16] ^ (uint16_t) invert_xor;
            x += x_step;
        }
        y += y_step;
    }
}

/* ---------------------------------------------------------------------------------------- */
/* imageconvert: src/modules/core/filter_imageconvert.c:58-419, macros mlt_frame.h:206-222   */

static inline void orc_yuv2rgb(int y, int u, int v, int *r, int *g, int *b)
{
    /* YUV2RGB_601_SCALED mlt_frame.h:216-222; >> on negatives is arithmetic */
    *r = (1192 * (y - 16) + 1634 * (v - 128)) >> 10;
    *g = (1192 * (y - 16) - 832 * (v - 128) - 401 * (u - 128)) >> 10;
    *b = (1192 * (y - 16) + 2066 * (u - 128)) >> 10;
    *r = *r < 0 ? 0 : *r > 255 ? 255 : *r;
    *g = *g < 0 ? 0 : *g > 255 ? 255 : *g;
    *b = *b < 0 ? 0 : *b > 255 ? 255 : *b;
}

static inline void orc_rgb2yuv(int r, int g, int b, int *y, int *u, int *v)
{
    /* RGB2YUV_601_SCALED mlt_frame.h:206-209; no clamp */
    *y = ((263 * r + 516 * g + 100 * b) >> 10) + 16;
    *u = ((-152 * r - 300 * g + 450 * b) >> 10) + 128;
    *v = ((450 * r - 377 * g - 73 * b) >> 10) + 128;
}

enum { F_RGB = 1, F_RGBA = 2, F_YUV422 = 3, F_YUV420P = 4, F_RGBA64 = 10 };


/* Returns 0 if handled, 1 if the reference's matrix cell is NULL (conversion_matrix :423-437)
 * and this project defines nothing for it.  Strides tightly packed (mlt_image_set_values).
 * dst bytes the reference leaves untouched (odd last column of yuv422->rgb(a)) stay untouched. */
ORC_API int orc_convert(int sf, int df, int w, int h, const uint8_t *src, uint8_t *dst,
                        const uint8_t *src_alpha, uint8_t *dst_alpha)
{
    int r, g, b, y0, y1, u0, u1, v0, v1;
    if (sf == F_YUV422 && (df == F_RGBA || df == F_RGB)) { /* :58-142 */
        int n = df == F_RGBA ? 4 : 3;
        for (int line = 0; line < h; line++) {
            const uint8_t *s = src + (size_t) line * w * 2;
            const uint8_t *a = src_alpha ? src_alpha + (size_t) line * w : NULL;
            uint8_t *d = dst + (size_t) line * w * n;
            for (int p = 0; p < w / 2; p++) {
                for (int k = 0; k < 2; k++) {
                    orc_yuv2rgb(s[2 * k], s[1], s[3], &r, &g, &b);
                    d[0] = r, d[1] = g, d[2] = b;
                    if (n == 4)
                        d[3] = a ? *a++ : 0xff;
                    d += n;
                }
                s += 4;
            }
        }
        return 0;
    }
    if ((sf == F_RGBA || sf == F_RGB) && df == F_YUV422) { /* :144-222 */
        int n = sf == F_RGBA ? 4 : 3;
        for (int line = 0; line < h; line++) {
            const uint8_t *s = src + (size_t) line * w * n;
            uint8_t *d = dst + (size_t) line * w * 2;
            uint8_t *a = (n == 4 && dst_alpha) ? dst_alpha + (size_t) line * w : NULL;
            for (int p = 0; p < w / 2; p++) {
                orc_rgb2yuv(s[0], s[1], s[2], &y0, &u0, &v0);
                orc_rgb2yuv(s[n], s[n + 1], s[n + 2], &y1, &u1, &v1);
                if (a)
                    *a++ = s[3], *a++ = s[7];
                d[0] = y0, d[1] = (u0 + u1) >> 1, d[2] = y1, d[3] = (v0 + v1) >> 1;
                s += 2 * n, d += 4;
            }
            if (w % 2) {
                orc_rgb2yuv(s[0], s[1], s[2], &y0, &u0, &v0);
                if (a)
                    *a++ = s[3];
                d[0] = y0, d[1] = u0;
            }
        }
        return 0;
    }
    if (sf == F_YUV420P && (df == F_YUV422 || df == F_RGB || df == F_RGBA)) { /* :224-325 */
        /* mlt_image_set_values: planes Y (w*h), U, V (w/2 * h/2), strides w, w/2, w/2;
         * the reference addresses chroma with strides[1]*line/2 (byte offset, not row). */
        const uint8_t *py = src, *pu = src + (size_t) w * h, *pv = pu + (size_t) (w / 2) * (h / 2);
        int n = df == F_YUV422 ? 0 : df == F_RGB ? 3 : 4;
        for (int line = 0; line < h; line++) {
            const uint8_t *sy = py + (size_t) w * line;
            const uint8_t *su = pu + (w / 2) * line / 2;
            const uint8_t *sv = pv + (w / 2) * line / 2;
            const uint8_t *a = (n == 4 && src_alpha) ? src_alpha + (size_t) line * w : NULL;
            uint8_t *d = dst + (size_t) line * w * (n ? n : 2);
            for (int p = 0; p < w / 2; p++) {
                if (n == 0) {
                    d[0] = sy[0], d[1] = *su, d[2] = sy[1], d[3] = *sv;
                    d += 4;
                } else {
                    for (int k = 0; k < 2; k++) {
                        orc_yuv2rgb(sy[k], *su, *sv, &r, &g, &b);
                        d[0] = r, d[1] = g, d[2] = b;
                        if (n == 4)
                            d[3] = a ? *a++ : 0xff;
                        d += n;
                    }
                }
                sy += 2, su++, sv++;
            }
        }
        return 0;
    }
    if (sf == F_YUV422 && df == F_YUV420P) { /* :327-368 */
        uint8_t *py = dst, *pu = dst + (size_t) w * h, *pv = pu + (size_t) (w / 2) * (h / 2);
        for (int line = 0; line < h; line++)
            for (int p = 0; p < w; p++)
                py[(size_t) line * w + p] = src[(size_t) line * w * 2 + 2 * p];
        for (int line = 0; line < h / 2; line++)
            for (int p = 0; p < w / 2; p++) {
                pu[(size_t) line * (w / 2) + p] = src[(size_t) line * 2 * w * 2 + 4 * p + 1];
                pv[(size_t) line * (w / 2) + p] = src[(size_t) line * 2 * w * 2 + 4 * p + 3];
            }
        return 0;
    }
    if (sf == F_RGB && df == F_RGBA) { /* :370-397 */
        for (size_t i = 0; i < (size_t) w * h; i++) {
            dst[4 * i] = src[3 * i], dst[4 * i + 1] = src[3 * i + 1], dst[4 * i + 2] = src[3 * i + 2];
            dst[4 * i + 3] = src_alpha ? src_alpha[i] : 0xff;
        }
        return 0;
    }
    if (sf == F_RGBA && df == F_RGB) { /* :399-419 */
        for (size_t i = 0; i < (size_t) w * h; i++) {
            dst[3 * i] = src[4 * i], dst[3 * i + 1] = src[4 * i + 1], dst[3 * i + 2] = src[4 * i + 2];
            if (dst_alpha)
                dst_alpha[i] = src[4 * i + 3];
        }
        return 0;
    }
    /* ---- project-defined rgba64 cells ("parity unpinned", see header of include/mltcuda.h) */
    if (sf == F_RGBA && df == F_RGBA64) {
        uint16_t *d = (uint16_t *) dst;
        for (size_t i = 0; i < (size_t) w * h * 4; i++)
            d[i] = (uint16_t) (src[i] * 257);
        return 0;
    }
    if (sf == F_RGBA64 && df == F_RGBA) {
        const uint16_t *s = (const uint16_t *) src;
        for (size_t i = 0; i < (size_t) w * h * 4; i++)
            dst[i] = (uint8_t) (s[i] >> 8);
        return 0;
    }
    if (sf == F_YUV422 && df == F_RGBA64) {
        uint16_t *d16 = (uint16_t *) dst;
        for (int line = 0; line < h; line++) {
            const uint8_t *s = src + (size_t) line * w * 2;
            const uint8_t *a = src_alpha ? src_alpha + (size_t) line * w : NULL;
            uint16_t *d = d16 + (size_t) line * w * 4;
            for (int p = 0; p < w / 2; p++) {
                for (int k = 0; k < 2; k++) {
                    orc_yuv2rgb(s[2 * k], s[1], s[3], &r, &g, &b);
                    d[0] = r * 257, d[1] = g * 257, d[2] = b * 257;
                    d[3] = a ? (uint16_t) (*a++ << 8) : 0xffff; /* filter_avcolour_space.c:337-349 */
                    d += 4;
                }
                s += 4;
            }
        }
        return 0;
    }
    if (sf == F_RGBA64 && df == F_YUV422) {
        const uint16_t *s16 = (const uint16_t *) src;
        for (int line = 0; line < h; line++) {
            const uint16_t *s = s16 + (size_t) line * w * 4;
            uint8_t *d = dst + (size_t) line * w * 2;
            uint8_t *a = dst_alpha ? dst_alpha + (size_t) line * w : NULL;
            for (int p = 0; p < w / 2; p++) {
                orc_rgb2yuv(s[0] >> 8, s[1] >> 8, s[2] >> 8, &y0, &u0, &v0);
                orc_rgb2yuv(s[4] >> 8, s[5] >> 8, s[6] >> 8, &y1, &u1, &v1);
                if (a) { /* filter_avcolour_space.c:257-270: lrint((float)a16/256.0) */
                    *a++ = (uint8_t) lrint((float) s[3] / 256.0);
                    *a++ = (uint8_t) lrint((float) s[7] / 256.0);
                }
                d[0] = y0, d[1] = (u0 + u1) >> 1, d[2] = y1, d[3] = (v0 + v1) >> 1;
                s += 8, d += 4;
            }
            if (w % 2) {
                orc_rgb2yuv(s[0] >> 8, s[1] >> 8, s[2] >> 8, &y0, &u0, &v0);
                if (a)
                    *a++ = (uint8_t) lrint((float) s[3] / 256.0);
                d[0] = y0, d[1] = u0;
            }
        }
        return 0;
    }
    return 1;
}

/* ---------------------------------------------------------------------------------------- */
/* luma: src/modules/core/transition_luma.c                                                  */

static inline float orc_smoothstep_f(float e1, float e2, float a) /* :310-321 */
{
    if (a < e1)
        return 0.f;
    if (a >= e2)
        return 1.f;
    a = (a - e1) / (e2 - e1);
    return (a * a) * (3 - (2 * a));
}

static inline int orc_all_ff(const uint8_t *p, size_t n) /* is_opaque :34-41 */
{
    for (size_t i = 0; i < n; i++)
        if (p[i] != 0xff)
            return 0;
    return 1;
}

#define ORC_CLAMP(x, lo, hi) ((x) < (lo) ? (lo) : (x) > (hi) ? (hi) : (x))

/* luma_composite_yuv422 (:327-473) after the images have been fetched.
 * translucent: -1 => decide like the reference (:372-373). Returns the flag used. */
ORC_API int orc_luma_yuv422(uint8_t *p_dest, const uint8_t *p_src, uint8_t *alpha_dest,
                            uint8_t *alpha_src, const uint16_t *map, int luma_width,
                            int luma_height, int width, int height, int width_dest,
                            int height_dest, int width_src, int height_src, float pos,
                            float frame_delta, float softness, int field_order, int invert,
                            int alpha_over, int translucent)
{
    (void) luma_height;
    if (width == 0 || height == 0)
        return 0;
    int is_translucent = translucent;
    if (translucent < 0)
        is_translucent = (alpha_dest && !orc_all_ff(alpha_dest, (size_t) width_dest * height_dest))
                         || (alpha_src && !orc_all_ff(alpha_src, (size_t) width_src * height_src));
    width_src = width_src > width_dest ? width_dest : width_src;
    height_src = height_src > height_dest ? height_dest : height_src;
    int stride_src = width_src * 2, stride_dest = width_dest * 2;
    float field_pos[2];
    field_pos[0] = (pos + ((field_order == 0 ? 1 : 0) * frame_delta * 0.5f)) * (1.f + softness);
    field_pos[1] = (pos + ((field_order == 0 ? 0 : 1) * frame_delta * 0.5f)) * (1.f + softness);
    int32_t x_diff = (luma_width << 16) / width, y_diff = (luma_height << 16) / height;
    uint32_t i_softness = (uint32_t) (softness * (1 << 16));
    int field_count = field_order < 0 ? 1 : 2;
    for (int field = 0; field < field_count; field++) {
        const uint8_t *p_row = p_src + field * stride_src;
        uint8_t *q_row = p_dest + field * stride_dest;
        int32_t y_offset = field << 16;
        for (int i = field; i < height_src; i += field_count) {
            const uint8_t *p = p_row;
            uint8_t *q = q_row;
            const uint16_t *l = map + (y_offset >> 16) * (luma_width * field_count);
            int32_t x_offset = 0;
            for (int j = 0; j < width_src; j++) {
                if (is_translucent) { /* :417-451 */
                    float weight = l[x_offset >> 16] / 65535.f;
                    float value = orc_smoothstep_f(weight, softness + weight, field_pos[field]);
                    float mix_a = (1.0f - value) * (alpha_dest ? *alpha_dest : 255) / 255.f;
                    float mix_b = value * (alpha_src ? *alpha_src : 255) / 255.f;
                    if (invert && alpha_src) {
                        float mix2 = mix_b + mix_a;
                        if (alpha_over) {
                            *alpha_src = (uint8_t) (255 * mix2);
                            if (mix2 != 0.f)
                                mix_b /= mix2;
                        } else {
                            float t = *alpha_src + 255 * mix_b;
                            *alpha_src = (uint8_t) ORC_CLAMP(t, 0, 255);
                        }
                    } else if (!invert && alpha_dest) {
                        float mix2 = mix_b + mix_a;
                        if (alpha_over) {
                            *alpha_dest = (uint8_t) (255 * mix2);
                            if (mix2 != 0.f)
                                mix_b /= mix2;
                        } else {
                            float t = *alpha_dest + 255 * mix_b;
                            *alpha_dest = (uint8_t) ORC_CLAMP(t, 0, 255);
                        }
                    }
                    q[0] = (uint8_t) (p[0] * mix_b + q[0] * (1.f - mix_b));
                    q[1] = (uint8_t) (p[1] * mix_b + q[1] * (1.f - mix_b));
                    if (alpha_dest)
                        alpha_dest++; /* NB: running pointers, not per-row (:446-449) */
                    if (alpha_src)
                        alpha_src++;
                } else { /* :452-463 */
                    uint16_t weight = l[x_offset >> 16];
                    uint32_t value = (uint32_t) orc_smoothstep(weight, (int32_t) (i_softness + weight),
                                                               (uint32_t) ((1 << 16) * field_pos[field]));
                    q[0] = (uint8_t) ((p[0] * value + q[0] * ((1 << 16) - value)) >> 16);
                    q[1] = (uint8_t) ((p[1] * value + q[1] * ((1 << 16) - value)) >> 16);
                }
                p += 2, q += 2;
                x_offset += x_diff;
            }
            y_offset += y_diff;
            p_row += field_count * stride_src;
            q_row += field_count * stride_dest;
        }
    }
    return is_translucent;
}

/* luma_composite_rgba32 / _rgba64 (:475-597, :599-723). bits = 8 or 16. Both images w x h. */
ORC_API int orc_luma_rgba(int bits, void *dest, const void *src, const uint16_t *map,
                          int luma_width, int luma_height, int w, int h, float pos,
                          float frame_delta, float softness, int field_order, int invert,
                          int alpha_over, int translucent)
{
    (void) invert; /* both branches of :551-567 are identical */
    int is_translucent = translucent;
    const float maxv = bits == 8 ? 255.f : 65535.f;
    if (translucent < 0) { /* mlt_image_is_opaque on both */
        is_translucent = 0;
        for (size_t i = 0; i < (size_t) w * h && !is_translucent; i++) {
            if (bits == 8)
                is_translucent = ((uint8_t *) dest)[4 * i + 3] != 0xff
                                 || ((const uint8_t *) src)[4 * i + 3] != 0xff;
            else
                is_translucent = ((uint16_t *) dest)[4 * i + 3] != 0xffff
                                 || ((const uint16_t *) src)[4 * i + 3] != 0xffff;
        }
    }
    float field_pos[2];
    field_pos[0] = (pos + ((field_order == 0 ? 1 : 0) * frame_delta * 0.5f)) * (1.f + softness);
    field_pos[1] = (pos + ((field_order == 0 ? 0 : 1) * frame_delta * 0.5f)) * (1.f + softness);
    int32_t x_diff = (luma_width << 16) / w, y_diff = (luma_height << 16) / h;
    uint32_t i_softness = (uint32_t) (softness * (1 << 16));
    int field_count = field_order < 0 ? 1 : 2;
    for (int field = 0; field < field_count; field++) {
        int32_t y_offset = field << 16;
        for (int i = field; i < h; i += field_count) {
            const uint16_t *l = map + (y_offset >> 16) * (luma_width * field_count);
            int32_t x_offset = 0;
            for (int j = 0; j < w; j++) {
                size_t o = ((size_t) i * w + j) * 4;
                uint32_t pv[4], qv[4];
                for (int c = 0; c < 4; c++) {
                    pv[c] = bits == 8 ? ((const uint8_t *) src)[o + c] : ((const uint16_t *) src)[o + c];
                    qv[c] = bits == 8 ? ((uint8_t *) dest)[o + c] : ((uint16_t *) dest)[o + c];
                }
                if (is_translucent) {
                    float weight = l[x_offset >> 16] / 65535.f;
                    float value = orc_smoothstep_f(weight, softness + weight, field_pos[field]);
                    float mix_a = (1.0f - value) * qv[3] / maxv;
                    float mix_b = value * pv[3] / maxv;
                    float mix2 = mix_b + mix_a;
                    float na;
                    if (alpha_over) {
                        na = maxv * mix2;
                        if (mix2 != 0.f)
                            mix_b /= mix2;
                    } else {
                        na = qv[3] + maxv * mix_b;
                        na = ORC_CLAMP(na, 0, maxv);
                    }
                    qv[3] = bits == 8 ? (uint8_t) na : (uint16_t) na;
                    for (int c = 0; c < 3; c++) {
                        float t = pv[c] * mix_b + qv[c] * (1.f - mix_b);
                        qv[c] = bits == 8 ? (uint8_t) t : (uint16_t) t;
                    }
                } else {
                    uint16_t weight = l[x_offset >> 16];
                    uint32_t value = (uint32_t) orc_smoothstep(weight, (int32_t) (i_softness + weight),
                                                               (uint32_t) ((1 << 16) * field_pos[field]));
                    for (int c = 0; c < 3; c++)
                        qv[c] = (pv[c] * value + qv[c] * ((1u << 16) - value)) >> 16;
                }
                for (int c = 0; c < 4; c++) {
                    if (bits == 8)
                        ((uint8_t *) dest)[o + c] = (uint8_t) qv[c];
                    else
                        ((uint16_t *) dest)[o + c] = (uint16_t) qv[c];
                }
                x_offset += x_diff;
            }
            y_offset += y_diff;
        }
    }
    return is_translucent;
}

/* composite_line_yuv_float rows (dissolve_slice_yuv422 :63-129) */
ORC_API void orc_dissolve_yuv422_float(uint8_t *dest, const uint8_t *src, uint8_t *alpha_a,
                                       const uint8_t *alpha_b, int width, int height, float weight)
{
    for (size_t i = 0; i < (size_t) width * height; i++) {
        float mix_a = (1.0f - weight) * (alpha_a ? alpha_a[i] : 255) / 255.f;
        float mix_b = weight * (alpha_b ? alpha_b[i] : 255) / 255.f;
        if (alpha_a) {
            float mix2 = mix_b + mix_a;
            alpha_a[i] = (uint8_t) (255 * mix2);
            if (mix2 != 0.f)
                mix_b /= mix2;
        }
        dest[2 * i] = (uint8_t) (src[2 * i] * mix_b + dest[2 * i] * (1.f - mix_b));
        dest[2 * i + 1] = (uint8_t) (src[2 * i + 1] * mix_b + dest[2 * i + 1] * (1.f - mix_b));
    }
}

/* dissolve_slice_rgba32/_rgba64 (:192-214, :243-265) */
ORC_API void orc_dissolve_rgba(int bits, void *dest, const void *src, int width, int height,
                               float weight)
{
    const float maxv = bits == 8 ? 255.f : 65535.f;
    for (size_t i = 0; i < (size_t) width * height; i++) {
        uint32_t pv[4], qv[4];
        for (int c = 0; c < 4; c++) {
            pv[c] = bits == 8 ? ((const uint8_t *) src)[4 * i + c] : ((const uint16_t *) src)[4 * i + c];
            qv[c] = bits == 8 ? ((uint8_t *) dest)[4 * i + c] : ((uint16_t *) dest)[4 * i + c];
        }
        float mix_a = (1.0f - weight) * qv[3] / maxv;
        float mix_b = weight * pv[3] / maxv;
        float mix2 = mix_b + mix_a;
        if (mix2 != 0.f)
            mix_b /= mix2;
        for (int c = 0; c < 3; c++) {
            float t = pv[c] * mix_b + qv[c] * (1.f - mix_b);
            qv[c] = bits == 8 ? (uint8_t) t : (uint16_t) t;
        }
        float na = maxv * mix2;
        qv[3] = bits == 8 ? (uint8_t) na : (uint16_t) na;
        for (int c = 0; c < 4; c++) {
            if (bits == 8)
                ((uint8_t *) dest)[4 * i + c] = (uint8_t) qv[c];
            else
                ((uint16_t *) dest)[4 * i + c] = (uint16_t) qv[c];
        }
    }
}

/* ---------------------------------------------------------------------------------------- */
/* affine: src/modules/plus/transition_affine.c:213-227,342-409 + src/modules/plus/interp.h  */

/* x86-64 C float->narrow-unsigned store: cvttss2si to int32 then truncation; NaN -> INT_MIN -> 0 */
static inline uint32_t orc_f2u(float f, uint32_t mask)
{
    int32_t i;
    if (!(f == f) || f >= 2147483648.f || f < -2147483648.f)
        i = (int32_t) 0x80000000;
    else
        i = (int32_t) f;
    return (uint32_t) i & mask;
}

/* Operation order switch.  0 = the C source read literally (strict IEEE, no reassociation).
 * 1 = the order the reference's Release build actually executes (gcc -O3 -ffast-math on x86-64,
 * read off the disassembly of oracle/_ref/lib/mlt/libmltplus.so): `a / maxv * o` is evaluated as
 * a * (o * (1/maxv)), `d / maxv` as d * (1/maxv), `/ j` as * (1/j), and sliced_proc hoists
 * (m01*y + m02) per row and multiplies by 1/dz.  tests/test_oracle_pinning.py checks mode 1
 * against the compiled interp*_b32/_b64 with ZERO mismatches for NN and BL. */
static int g_compiled_order = 0;
ORC_API void orc_set_compiled_order(int on)
{
    g_compiled_order = on;
}

/* one sample+blend; bits 8 or 16; kind 0 NN (:50-86), 1 BL (:90-194), 2 BC (:198-334) */
static void orc_interp(int bits, int kind, const void *sv, int w, int h, float x, float y, float o,
                       void *dv, int is_atop)
{
    const float maxv = bits == 8 ? 255.0f : 65535.0f;
    const uint32_t mask = bits == 8 ? 0xff : 0xffff;
#define S(i) (bits == 8 ? (int) ((const uint8_t *) sv)[i] : (int) ((const uint16_t *) sv)[i])
#define D(i) (bits == 8 ? (int) ((uint8_t *) dv)[i] : (int) ((uint16_t *) dv)[i])
#define SETD(i, val) \
    do { \
        if (bits == 8) \
            ((uint8_t *) dv)[i] = (uint8_t) (val); \
        else \
            ((uint16_t *) dv)[i] = (uint16_t) (val); \
    } while (0)
    if (kind == 0) {
        int p = (int) rintf(x) * 4 + (int) rintf(y) * 4 * w;
        float alpha_s = (float) S(p + 3) / maxv * o;
        float alpha_d = (float) D(3) / maxv;
        if (g_compiled_order) {
            alpha_s = (float) S(p + 3) * (o * (1.0f / maxv));
            alpha_d = (float) D(3) * (1.0f / maxv);
        }
        float alpha = alpha_s + alpha_d - alpha_s * alpha_d;
        if (is_atop)
            SETD(3, S(p + 3));
        else
            SETD(3, orc_f2u(maxv * alpha, mask));
        alpha = alpha_s / alpha;
        for (int c = 0; c < 3; c++)
            SETD(c, orc_f2u(D(c) * (1.0f - alpha) + S(p + c) * alpha, mask));
    } else if (kind == 1) {
        int m = (int) floorf(x);
        if (m + 2 > w)
            m = w - 2;
        int n = (int) floorf(y);
        if (n + 2 > h)
            n = h - 2;
        int k = 4 * (n * w + m), l = 4 * ((n + 1) * w + m);
        int k1 = k + 4, l1 = l + 4;
        float fx = x - (float) m, fy = y - (float) n;
        float a = S(k + 3) + (S(k1 + 3) - S(k + 3)) * fx;
        float b = S(l + 3) + (S(l1 + 3) - S(l + 3)) * fx;
        float alpha_s = a + (b - a) * fy;
        float alpha_d = g_compiled_order ? (float) D(3) * (1.0f / maxv) : (float) D(3) / maxv;
        if (is_atop)
            SETD(3, orc_f2u(alpha_s, mask));
        alpha_s = g_compiled_order ? alpha_s * (o * (1.0f / maxv)) : alpha_s / maxv * o;
        float alpha = alpha_s + alpha_d - alpha_s * alpha_d;
        if (!is_atop)
            SETD(3, orc_f2u(maxv * alpha, mask));
        alpha = alpha_s / alpha;
        for (int c = 0; c < 3; c++) {
            a = S(k + c) + (S(k1 + c) - S(k + c)) * fx;
            b = S(l + c) + (S(l1 + c) - S(l + c)) * fx;
            SETD(c, orc_f2u(D(c) * (1.0f - alpha) + (a + (b - a) * fy) * alpha, mask));
        }
    } else {
        int m = (int) ceilf(x) - 2;
        if (m < 0)
            m = 0;
        if (m + 5 > w)
            m = w - 4;
        int n = (int) ceilf(y) - 2;
        if (n < 0)
            n = 0;
        if (n + 5 > h)
            n = h - 4;
        float alpha = 1.0f;
        for (int b = 3; b > -1; b--) {
            float p[4], pc[4][4];
            for (int i = 0; i < 4; i++) {
                int l = m + (i + n) * w;
                for (int c = 0; c < 4; c++)
                    pc[c][i] = (float) S(4 * (l + c) + b);
            }
            for (int j = 1; j < 4; j++)
                for (int i = 3; i >= j; i--) {
                    float kk = g_compiled_order ? (y - i - n) * (1.0f / j) : (y - i - n) / j;
                    for (int c = 0; c < 4; c++)
                        pc[c][i] = pc[c][i] + kk * (pc[c][i] - pc[c][i - 1]);
                }
            for (int c = 0; c < 4; c++)
                p[c] = pc[c][3];
            for (int j = 1; j < 4; j++)
                for (int i = 3; i >= j; i--)
                    p[i] = p[i] + (g_compiled_order ? (x - i - m) * (1.0f / j) : (x - i - m) / j) * (p[i] - p[i - 1]);
            if (p[3] < 0.0f)
                p[3] = 0.0f;
            if (p[3] > maxv)
                p[3] = maxv;
            if (b == 3) {
                float alpha_s = g_compiled_order ? p[3] * (o * (1.0f / maxv)) : p[3] / maxv * o;
                float alpha_d = g_compiled_order ? (float) D(3) * (1.0f / maxv) : (float) D(3) / maxv;
                alpha = alpha_s + alpha_d - alpha_s * alpha_d;
                SETD(3, is_atop ? orc_f2u(p[3], mask) : orc_f2u(maxv * alpha, mask));
                alpha = alpha_s / alpha;
            } else {
                SETD(b, orc_f2u(D(b) * (1.0f - alpha) + p[3] * alpha, mask));
            }
        }
    }
#undef S
#undef D
#undef SETD
}

ORC_API void orc_interp_px(int bits, int kind, const void *s, int w, int h, float x, float y,
                           float o, void *d, int is_atop)
{
    orc_interp(bits, kind, s, w, h, x, y, o, d, is_atop);
}

/* sliced_proc (:342-409), all rows. matrix is row-major double[9]. Returns covered pixel count. */
ORC_API long orc_affine(int bits, void *a_image, const void *b_image, int a_width, int a_height,
                        int b_width, int b_height, const double *m, double lower_x, double lower_y,
                        double dz, double mix, double x_offset, double y_offset, double minima,
                        double xmax, double ymax, int b_alpha, int interp)
{
    long covered = 0;
    size_t px = bits == 8 ? 4 : 8;
    double y = lower_y;
    const double rdz = 1.0 / dz;
    for (int i = 0; i < a_height; i++, y++) {
        double x = lower_x;
        const double rowx = m[1] * y + m[2], rowy = m[4] * y + m[5];
        for (int j = 0; j < a_width; j++, x++) {
            double dx = (m[0] * x + m[1] * y + m[2]) / dz + x_offset;
            double dy = (m[3] * x + m[4] * y + m[5]) / dz + y_offset;
            if (g_compiled_order) {
                dx = (m[0] * x + rowx) * rdz + x_offset;
                dy = (m[3] * x + rowy) * rdz + y_offset;
            }
            if (dx >= minima && dx <= xmax && dy >= minima && dy <= ymax) {
                orc_interp(bits, interp, b_image, b_width, b_height, (float) dx, (float) dy,
                           (float) mix, (uint8_t *) a_image + ((size_t) i * a_width + j) * px,
                           b_alpha);
                covered++;
            }
        }
    }
    return covered;
}

/* ---------------------------------------------------------------------------------------- */
/* producers / normalisers (SURVEY 8 f1,f2)                                                  */

/* producer_noise.c:37-51 (MWC RNG), :108-123 (fill from the end, seed from position) */
ORC_API void orc_noise(uint8_t *image, int width, int height, int position)
{
    uint32_t x = 521288629u + (uint32_t) position - ((uint32_t) position << 16);
    uint32_t y = 362436069u - (uint32_t) position + ((uint32_t) position << 16);
    uint8_t *p = image + (size_t) width * height * 2;
    while (p != image) {
        x = 18000 * (x & 65535) + (x >> 16);
        y = 30903 * (y & 65535) + (y >> 16);
        uint32_t value = ((x << 16) + (y & 65535)) & 0xff;
        *(--p) = 128;
        *(--p) = value < 16 ? 16 : value > 240 ? 240 : (uint8_t) value;
    }
}

/* filter_rescale.c:48-120 nearest-neighbour yuv422 scaling (filter_scale) */
ORC_API void orc_rescale_yuv422(uint8_t *dst, int ow, int oh, const uint8_t *src, int iw, int ih)
{
    int istride = iw * 2, ostride = ow * 2;
    iw -= iw % 4;
    int out_x_range = ow / 2, out_y_range = oh / 2, in_x_range = iw / 2, in_y_range = ih / 2;
    const uint8_t *in_middle = src + istride * in_y_range + in_x_range * 2;
    int scale_width = (iw << 16) / ow, scale_height = (ih << 16) / oh;
    int outer = out_x_range * scale_width, bottom = out_y_range * scale_height;
    uint8_t *out_line = dst;
    for (int dy = -bottom; dy < bottom; dy += scale_height) {
        uint8_t *o = out_line;
        const uint8_t *in_line = in_middle + (dy >> 16) * istride;
        for (int dx = -outer; dx < outer; dx += scale_width) {
            int base = dx >> 15;
            base &= 0xfffffffe;
            *o++ = in_line[base];
            base &= 0xfffffffc;
            *o++ = in_line[base + 1];
            dx += scale_width;
            base = dx >> 15;
            base &= 0xfffffffe;
            *o++ = in_line[base];
            base &= 0xfffffffc;
            *o++ = in_line[base + 3];
        }
        out_line += ostride;
    }
}

/* producer_colour.c:100-274: constant fill. yuv422 uses RGB2YUV_601_SCALED of the colour and an
 * alpha plane filled with a; rgba/rgba64 packed ((c<<8)+c for 16 bit, :196-205). */
ORC_API void orc_colour(void *dst, uint8_t *dst_alpha, int format, int width, int height,
                        uint8_t r, uint8_t g, uint8_t b, uint8_t a)
{
    size_t n = (size_t) width * height;
    if (format == F_YUV422) {
        int y, u, v;
        orc_rgb2yuv(r, g, b, &y, &u, &v);
        uint8_t *p = (uint8_t *) dst;
        for (int line = 0; line < height; line++) {
            for (int i = 0; i < width / 2; i++)
                *p++ = y, *p++ = u, *p++ = y, *p++ = v;
            if (width % 2)
                *p++ = y, *p++ = u;
        }
        if (dst_alpha)
            memset(dst_alpha, a, n);
    } else if (format == F_RGBA) {
        uint8_t *p = (uint8_t *) dst;
        for (size_t i = 0; i < n; i++)
            *p++ = r, *p++ = g, *p++ = b, *p++ = a;
    } else if (format == F_RGBA64) {
        uint16_t *p = (uint16_t *) dst;
        for (size_t i = 0; i < n; i++)
            *p++ = (r << 8) + r, *p++ = (g << 8) + g, *p++ = (b << 8) + b, *p++ = (a << 8) + a;
    } else if (format == F_RGB) {
        uint8_t *p = (uint8_t *) dst;
        for (size_t i = 0; i < n; i++)
            *p++ = r, *p++ = g, *p++ = b;
    }
}

/* FNV-1a 64 over a buffer, as oracle/mlt_dump.c hashes the reference's frames */
ORC_API uint64_t orc_fnv1a64(const uint8_t *p, size_t n)
{
    uint64_t h = 1469598103934665603ULL;
    for (size_t i = 0; i < n; i++) {
        h ^= p[i];
        h *= 1099511628211ULL;
    }
    return h;
}
