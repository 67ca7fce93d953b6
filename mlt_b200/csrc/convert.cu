// convert.cu -- K1: imageconvert cells (reference src/modules/core/filter_imageconvert.c:58-437).
//
// Two tiers, both bit-exact:
//  * vector kernels for the hot cells (yuv422<->rgba, yuv422<->rgba64, rgba<->rgba64): one thread
//    = 8 pixels = one 16-byte YUYV word group, 128-bit coalesced loads/stores;
//  * a generic pixel-pair kernel for every other cell and for odd widths / unaligned strides.
// HBM-bound streaming: algorithmic bytes 6 B/px (yuv422->rgba, +1 with alpha plane), 7 B/px
// (rgba->yuv422 + alpha plane).
#include "common.cuh"
#include "pixel_ops.cuh"

namespace {

struct CvtParams
{
    const uint8_t *src;
    uint8_t *dst;
    const uint8_t *src_alpha;
    uint8_t *dst_alpha;
    int width, height;       // logical image
    int rows, row_px;        // iteration space of the vector kernels (linearised when tight)
    long src_stride, dst_stride; // bytes
    size_t src_bs, dst_bs, sa_bs, da_bs;
    int batch;
};

// ---- vector kernels -----------------------------------------------------------------------------

// yuv422 -> rgba : filter_imageconvert.c:58-111
__global__ void __launch_bounds__(256) k_yuv422_rgba_v(CvtParams P)
{
    const int cpr = (P.row_px + 7) >> 3;
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    long long total = (long long) cpr * P.rows * P.batch;
    if (gid >= total)
        return;
    int c = (int) (gid % cpr);
    long long t = gid / cpr;
    int r = (int) (t % P.rows);
    int b = (int) (t / P.rows);
    const uint8_t *s = P.src + b * P.src_bs + r * P.src_stride + (size_t) c * 16;
    uint8_t *d = P.dst + b * P.dst_bs + r * P.dst_stride + (size_t) c * 32;
    int px = P.row_px - c * 8;
    uint4 in = ldg16(s);
    uint32_t a[2] = {0xffffffffu, 0xffffffffu};
    if (P.src_alpha) {
        const uint8_t *ap = P.src_alpha + b * P.sa_bs + (size_t) r * P.row_px + (size_t) c * 8;
        if (px >= 8 && (((uintptr_t) ap) & 7) == 0) {
            uint2 av = ldg8(ap);
            a[0] = av.x, a[1] = av.y;
        } else {
            a[0] = a[1] = 0;
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k < px)
                    a[k >> 2] |= (uint32_t) __ldg(ap + k) << ((k & 3) * 8);
        }
    }
    uint32_t w[4] = {in.x, in.y, in.z, in.w};
    uint32_t o[8];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        mc_yuyv_to_rgba2w(w[k], a[k >> 1], (k & 1) * 2, (k & 1) * 2 + 1, o[2 * k], o[2 * k + 1]);
    }
    if (px >= 8) {
        reinterpret_cast<uint4 *>(d)[0] = make_uint4(o[0], o[1], o[2], o[3]);
        reinterpret_cast<uint4 *>(d)[1] = make_uint4(o[4], o[5], o[6], o[7]);
    } else {
        int pairs = px >> 1; // an odd trailing pixel is not written (:70 total = width/2 + 1)
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (k < pairs)
                reinterpret_cast<uint2 *>(d)[k] = make_uint2(o[2 * k], o[2 * k + 1]);
    }
}

// rgba -> yuv422 + alpha plane : filter_imageconvert.c:144-184 (even widths; odd -> generic)
__global__ void __launch_bounds__(256) k_rgba_yuv422_v(CvtParams P)
{
    const int cpr = (P.row_px + 7) >> 3;
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    long long total = (long long) cpr * P.rows * P.batch;
    if (gid >= total)
        return;
    int c = (int) (gid % cpr);
    long long t = gid / cpr;
    int r = (int) (t % P.rows);
    int b = (int) (t / P.rows);
    const uint8_t *s = P.src + b * P.src_bs + r * P.src_stride + (size_t) c * 32;
    uint8_t *d = P.dst + b * P.dst_bs + r * P.dst_stride + (size_t) c * 16;
    int px = P.row_px - c * 8;
    uint32_t p[8];
    if (px >= 8) {
        uint4 i0 = ldg16(s), i1 = ldg16(s + 16);
        p[0] = i0.x, p[1] = i0.y, p[2] = i0.z, p[3] = i0.w;
        p[4] = i1.x, p[5] = i1.y, p[6] = i1.z, p[7] = i1.w;
    } else {
#pragma unroll
        for (int k = 0; k < 8; k++)
            p[k] = k < px ? __ldg(reinterpret_cast<const uint32_t *>(s) + k) : 0;
    }
    uint32_t y[4], a[2];
#pragma unroll
    for (int k = 0; k < 4; k++)
        y[k] = mc_rgba2_to_yuyv(p[2 * k], p[2 * k + 1]);
    a[0] = (p[0] >> 24) | ((p[1] >> 24) << 8) | ((p[2] >> 24) << 16) | ((p[3] >> 24) << 24);
    a[1] = (p[4] >> 24) | ((p[5] >> 24) << 8) | ((p[6] >> 24) << 16) | ((p[7] >> 24) << 24);
    uint8_t *ap = P.dst_alpha ? P.dst_alpha + b * P.da_bs + (size_t) r * P.row_px + (size_t) c * 8 : nullptr;
    if (px >= 8) {
        *reinterpret_cast<uint4 *>(d) = make_uint4(y[0], y[1], y[2], y[3]);
        if (ap) {
            if ((((uintptr_t) ap) & 7) == 0)
                *reinterpret_cast<uint2 *>(ap) = make_uint2(a[0], a[1]);
            else {
#pragma unroll
                for (int k = 0; k < 8; k++)
                    ap[k] = (uint8_t) (a[k >> 2] >> ((k & 3) * 8));
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (k < (px >> 1))
                reinterpret_cast<uint32_t *>(d)[k] = y[k];
        if (ap) {
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k < px)
                    ap[k] = (uint8_t) (a[k >> 2] >> ((k & 3) * 8));
        }
    }
}

// ---- generic pixel-pair kernel --------------------------------------------------------------------

enum Cell {
    C_YUV422_RGB, C_YUV422_RGBA, C_RGB_YUV422, C_RGBA_YUV422, C_420_422, C_420_RGB, C_420_RGBA,
    C_RGB_RGBA, C_RGBA_RGB, C_RGBA_RGBA64, C_RGBA64_RGBA, C_YUV422_RGBA64, C_RGBA64_YUV422
};

__device__ __forceinline__ uint8_t lrint_div256(uint16_t a)
{
    // lrint((float) a / 256.0): round-half-even of an exact quotient (avformat/filter_avcolour_space.c:262)
    return (uint8_t) __double2int_rn((double) (float) a / 256.0);
}

template <int CELL> __global__ void __launch_bounds__(256) k_convert_pair(CvtParams P)
{
    const int w = P.width, h = P.height;
    const int ppr = (w + 1) >> 1; // pairs per row incl. a trailing single pixel
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    long long total = (long long) ppr * h * P.batch;
    if (gid >= total)
        return;
    int pi = (int) (gid % ppr);
    long long t = gid / ppr;
    int line = (int) (t % h);
    int b = (int) (t / h);
    const uint8_t *src = P.src + b * P.src_bs;
    uint8_t *dst = P.dst + b * P.dst_bs;
    const uint8_t *sa = P.src_alpha ? P.src_alpha + b * P.sa_bs + (size_t) line * w + 2 * pi : nullptr;
    uint8_t *da = P.dst_alpha ? P.dst_alpha + b * P.da_bs + (size_t) line * w + 2 * pi : nullptr;
    const int x = 2 * pi;
    const bool single = (x + 1 >= w); // odd trailing pixel

    if (CELL == C_YUV422_RGB || CELL == C_YUV422_RGBA || CELL == C_YUV422_RGBA64 || CELL == C_420_422
        || CELL == C_420_RGB || CELL == C_420_RGBA) {
        if (single)
            return; // reference loops width/2 pairs only
        uint32_t yw;
        if (CELL == C_420_422 || CELL == C_420_RGB || CELL == C_420_RGBA) {
            // :224-325; chroma row offset is (stride*line)/2 as written in the reference
            const uint8_t *py = src + (size_t) w * line + x;
            const uint8_t *pu = src + (size_t) w * h + ((w / 2) * line) / 2 + pi;
            const uint8_t *pv = src + (size_t) w * h + (size_t) (w / 2) * (h / 2) + ((w / 2) * line) / 2 + pi;
            yw = py[0] | (pu[0] << 8) | (py[1] << 16) | (pv[0] << 24);
        } else {
            // rows of odd-width yuv422 images are only 2-byte aligned
            const uint16_t *s16 = reinterpret_cast<const uint16_t *>(src + line * P.src_stride + 4 * (size_t) pi);
            yw = (uint32_t) s16[0] | ((uint32_t) s16[1] << 16);
        }
        if (CELL == C_420_422) {
            uint16_t *d16 = reinterpret_cast<uint16_t *>(dst + line * P.dst_stride + 4 * (size_t) pi);
            d16[0] = (uint16_t) yw, d16[1] = (uint16_t) (yw >> 16);
            return;
        }
        uint32_t a0 = sa ? sa[0] : 0xff, a1 = sa ? sa[1] : 0xff, p0, p1;
        mc_yuyv_to_rgba2(yw, a0, a1, p0, p1);
        if (CELL == C_YUV422_RGBA || CELL == C_420_RGBA) {
            uint32_t *d = reinterpret_cast<uint32_t *>(dst + line * P.dst_stride) + x;
            d[0] = p0, d[1] = p1;
        } else if (CELL == C_YUV422_RGB || CELL == C_420_RGB) {
            uint8_t *d = dst + line * P.dst_stride + 3 * (size_t) x;
            d[0] = p0, d[1] = p0 >> 8, d[2] = p0 >> 16, d[3] = p1, d[4] = p1 >> 8, d[5] = p1 >> 16;
        } else { // C_YUV422_RGBA64 (project-defined: v*257, alpha a8<<8 or 0xffff)
            uint16_t *d = reinterpret_cast<uint16_t *>(dst + line * P.dst_stride) + 4 * (size_t) x;
            uint32_t pp[2] = {p0, p1};
            for (int k = 0; k < 2; k++) {
                d[4 * k] = (pp[k] & 0xff) * 257, d[4 * k + 1] = ((pp[k] >> 8) & 0xff) * 257;
                d[4 * k + 2] = ((pp[k] >> 16) & 0xff) * 257;
                d[4 * k + 3] = sa ? (uint16_t) (sa[k] << 8) : 0xffff;
            }
        }
        return;
    }
    if (CELL == C_RGB_YUV422 || CELL == C_RGBA_YUV422 || CELL == C_RGBA64_YUV422) {
        uint32_t p[2] = {0, 0};
        int n = single ? 1 : 2;
        for (int k = 0; k < n; k++) {
            if (CELL == C_RGB_YUV422) {
                const uint8_t *s = src + line * P.src_stride + 3 * (size_t) (x + k);
                p[k] = s[0] | (s[1] << 8) | (s[2] << 16);
            } else if (CELL == C_RGBA_YUV422) {
                p[k] = reinterpret_cast<const uint32_t *>(src + line * P.src_stride)[x + k];
                if (da)
                    da[k] = p[k] >> 24;
            } else {
                const uint16_t *s = reinterpret_cast<const uint16_t *>(src + line * P.src_stride) + 4 * (size_t) (x + k);
                p[k] = (s[0] >> 8) | ((s[1] >> 8) << 8) | ((s[2] >> 8) << 16);
                if (da)
                    da[k] = lrint_div256(s[3]);
            }
        }
        uint8_t *d = dst + line * P.dst_stride + 2 * (size_t) x;
        if (single) { // :174-182
            int y0, u0, v0;
            mc_rgb2yuv(p[0] & 0xff, (p[0] >> 8) & 0xff, (p[0] >> 16) & 0xff, y0, u0, v0);
            d[0] = y0, d[1] = u0;
        } else {
            uint32_t yw = mc_rgba2_to_yuyv(p[0], p[1]);
            reinterpret_cast<uint16_t *>(d)[0] = (uint16_t) yw, reinterpret_cast<uint16_t *>(d)[1] = (uint16_t) (yw >> 16);
        }
        return;
    }
    // per-pixel cells
    int n = single ? 1 : 2;
    for (int k = 0; k < n; k++) {
        size_t xi = x + k;
        if (CELL == C_RGB_RGBA) { // :370-397
            const uint8_t *s = src + line * P.src_stride + 3 * xi;
            reinterpret_cast<uint32_t *>(dst + line * P.dst_stride)[xi]
                = s[0] | (s[1] << 8) | (s[2] << 16) | ((sa ? sa[k] : 0xffu) << 24);
        } else if (CELL == C_RGBA_RGB) { // :399-419
            uint32_t v = reinterpret_cast<const uint32_t *>(src + line * P.src_stride)[xi];
            uint8_t *d = dst + line * P.dst_stride + 3 * xi;
            d[0] = v, d[1] = v >> 8, d[2] = v >> 16;
            if (da)
                da[k] = v >> 24;
        } else if (CELL == C_RGBA_RGBA64) {
            uint32_t v = reinterpret_cast<const uint32_t *>(src + line * P.src_stride)[xi];
            uint2 o;
            o.x = ((v & 0xff) * 257) | ((((v >> 8) & 0xff) * 257) << 16);
            o.y = (((v >> 16) & 0xff) * 257) | (((v >> 24) * 257) << 16);
            reinterpret_cast<uint2 *>(dst + line * P.dst_stride)[xi] = o;
        } else if (CELL == C_RGBA64_RGBA) {
            uint2 v = reinterpret_cast<const uint2 *>(src + line * P.src_stride)[xi];
            reinterpret_cast<uint32_t *>(dst + line * P.dst_stride)[xi]
                = ((v.x >> 8) & 0xff) | (((v.x >> 24) & 0xff) << 8) | (((v.y >> 8) & 0xff) << 16) | ((v.y >> 24) << 24);
        }
    }
}

// ---- vector tier of the generic cells ----------------------------------------------------------------
// Same arithmetic as k_convert_pair, one thread = 8 pixels of a row held in registers: sources are read
// and results written as 8 / 16-byte vectors, bytes are moved with PRMT.  Needs width % 8 == 0 (16 for
// a yuv420p source, whose chroma rows start at ((w/2)*line)/2), 16-byte aligned images and row strides.
__device__ __forceinline__ uint32_t lrint_div256_i(uint32_t a)
{
    // lrint(a / 256.0) for a < 65536, round-half-even, without the conversion pipe; low byte kept by callers
    return (a + 127u + ((a >> 8) & 1u)) >> 8;
}

template <int CELL> __global__ void __launch_bounds__(256) k_convert_vec(CvtParams P)
{
    const int g = blockIdx.x * 32 + threadIdx.x, line = blockIdx.y * 8 + threadIdx.y;
    const int w = P.width, h = P.height;
    if (g >= (w >> 3) || line >= h)
        return;
    const int b = blockIdx.z;
    const uint8_t *src = P.src + b * P.src_bs;
    uint8_t *dst = P.dst + b * P.dst_bs;
    const size_t x = 8 * (size_t) g;
    constexpr bool SRC_422 = CELL == C_YUV422_RGB || CELL == C_YUV422_RGBA64;
    constexpr bool SRC_420 = CELL == C_420_422 || CELL == C_420_RGB || CELL == C_420_RGBA;
    constexpr bool SRC_RGB = CELL == C_RGB_YUV422 || CELL == C_RGB_RGBA;
    constexpr bool SRC_RGBA = CELL == C_RGBA_RGB || CELL == C_RGBA_RGBA64;
    constexpr bool SRC_RGBA64 = CELL == C_RGBA64_RGBA || CELL == C_RGBA64_YUV422;
    constexpr bool DST_RGB = CELL == C_YUV422_RGB || CELL == C_420_RGB || CELL == C_RGBA_RGB;
    constexpr bool DST_RGBA = CELL == C_420_RGBA || CELL == C_RGB_RGBA || CELL == C_RGBA64_RGBA;
    constexpr bool DST_422 = CELL == C_RGB_YUV422 || CELL == C_RGBA64_YUV422;
    constexpr bool DST_RGBA64 = CELL == C_YUV422_RGBA64 || CELL == C_RGBA_RGBA64;
    uint32_t yuv[4] = {0, 0, 0, 0}, px[8], al[2] = {0xffffffffu, 0xffffffffu}, a16[8];
    const bool has_sa = P.src_alpha != nullptr;
    if (has_sa && (SRC_422 || SRC_420 || SRC_RGB)) {
        const uint2 v = ldg8(P.src_alpha + b * P.sa_bs + (size_t) line * w + x);
        al[0] = v.x, al[1] = v.y;
    }
    if (SRC_422) {
        const uint4 v = ldg16(src + line * P.src_stride + 2 * x);
        yuv[0] = v.x, yuv[1] = v.y, yuv[2] = v.z, yuv[3] = v.w;
    } else if (SRC_420) { // :224-325; chroma row offset is (stride*line)/2 as written in the reference
        const uint2 yy = ldg8(src + (size_t) w * line + x);
        const size_t co = (size_t) (((w / 2) * line) / 2) + 4 * (size_t) g;
        const uint32_t uu = __ldg(reinterpret_cast<const uint32_t *>(src + (size_t) w * h + co));
        const uint32_t vv = __ldg(reinterpret_cast<const uint32_t *>(src + (size_t) w * h + (size_t) (w / 2) * (h / 2) + co));
        const uint32_t yw[2] = {yy.x, yy.y};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t t = __byte_perm(yw[k >> 1], uu, ((k & 1) ? 0x0302 : 0x0100) | ((4 + k) << 4)); // y, u, y', -
            yuv[k] = __byte_perm(t, vv, 0x0210 | ((4 + k) << 12));
        }
    } else if (SRC_RGB) {
        const uint8_t *sp = src + line * P.src_stride + 3 * x;
        const uint2 v0 = ldg8(sp), v1 = ldg8(sp + 8), v2 = ldg8(sp + 16);
        const uint32_t bw[6] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int i = (3 * k) >> 2, o = (3 * k) & 3, i1 = i + 1 > 5 ? 5 : i + 1;
            const uint32_t rgb = __byte_perm(bw[i], bw[i1], o | ((o + 1) << 4) | ((o + 2) << 8));
            px[k] = __byte_perm(rgb, al[k >> 2], 0x0210 | ((4 + (k & 3)) << 12));
        }
    } else if (SRC_RGBA) {
        const uint8_t *sp = src + line * P.src_stride + 4 * x;
        const uint4 v0 = ldg16(sp), v1 = ldg16(sp + 16);
        px[0] = v0.x, px[1] = v0.y, px[2] = v0.z, px[3] = v0.w, px[4] = v1.x, px[5] = v1.y, px[6] = v1.z, px[7] = v1.w;
    } else if (SRC_RGBA64) {
        const uint8_t *sp = src + line * P.src_stride + 8 * x;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint4 v = ldg16(sp + 16 * k);
            px[2 * k] = __byte_perm(v.x, v.y, 0x7531), px[2 * k + 1] = __byte_perm(v.z, v.w, 0x7531); // high bytes
            a16[2 * k] = v.y >> 16, a16[2 * k + 1] = v.w >> 16;
        }
    }
    if (SRC_422 || (SRC_420 && CELL != C_420_422)) {
#pragma unroll
        for (int k = 0; k < 4; k++)
            mc_yuyv_to_rgba2w(yuv[k], al[k >> 1], (k & 1) * 2, (k & 1) * 2 + 1, px[2 * k], px[2 * k + 1]);
    }
    if (CELL == C_420_422) {
        *reinterpret_cast<uint4 *>(dst + line * P.dst_stride + 2 * x) = make_uint4(yuv[0], yuv[1], yuv[2], yuv[3]);
    } else if (DST_RGBA) {
        uint4 *dp = reinterpret_cast<uint4 *>(dst + line * P.dst_stride + 4 * x);
        dp[0] = make_uint4(px[0], px[1], px[2], px[3]), dp[1] = make_uint4(px[4], px[5], px[6], px[7]);
    } else if (DST_RGB) {
        uint32_t ow[6];
#pragma unroll
        for (int m = 0; m < 6; m++) { // output bytes 4m .. 4m+3: pixel n/3, channel n%3
            uint32_t sel = 0;
            const int p0 = (4 * m) / 3;
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const int n = 4 * m + t, pi = n / 3, ch = n % 3;
                sel |= (uint32_t) ((pi - p0) * 4 + ch) << (4 * t);
            }
            ow[m] = __byte_perm(px[p0], px[p0 + 1 > 7 ? 7 : p0 + 1], sel);
        }
        uint2 *dp = reinterpret_cast<uint2 *>(dst + line * P.dst_stride + 3 * x);
        dp[0] = make_uint2(ow[0], ow[1]), dp[1] = make_uint2(ow[2], ow[3]), dp[2] = make_uint2(ow[4], ow[5]);
    } else if (DST_422) {
        uint32_t y[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            y[k] = mc_rgba2_to_yuyv(px[2 * k], px[2 * k + 1]);
        *reinterpret_cast<uint4 *>(dst + line * P.dst_stride + 2 * x) = make_uint4(y[0], y[1], y[2], y[3]);
    } else if (DST_RGBA64) { // v * 257 = (v << 8) | v; yuv422 source: alpha a8 << 8, or 0xffff without a plane
        const uint32_t hi_sel = (CELL == C_YUV422_RGBA64 && has_sa) ? 0x3422u : 0x3322u;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            reinterpret_cast<uint4 *>(dst + line * P.dst_stride + 8 * x)[k]
                = make_uint4(__byte_perm(px[2 * k], 0, 0x1100), __byte_perm(px[2 * k], 0, hi_sel),
                             __byte_perm(px[2 * k + 1], 0, 0x1100), __byte_perm(px[2 * k + 1], 0, hi_sel));
        }
    }
    if ((CELL == C_RGBA_RGB || CELL == C_RGBA64_YUV422) && P.dst_alpha) { // alpha byte -> new alpha plane
        uint32_t ao[2];
        if (CELL == C_RGBA_RGB) {
            ao[0] = __byte_perm(__byte_perm(px[0], px[1], 0x0073), __byte_perm(px[2], px[3], 0x0073), 0x5410);
            ao[1] = __byte_perm(__byte_perm(px[4], px[5], 0x0073), __byte_perm(px[6], px[7], 0x0073), 0x5410);
        } else {
            uint32_t r[8];
#pragma unroll
            for (int k = 0; k < 8; k++)
                r[k] = lrint_div256_i(a16[k]);
            ao[0] = __byte_perm(__byte_perm(r[0], r[1], 0x0040), __byte_perm(r[2], r[3], 0x0040), 0x5410);
            ao[1] = __byte_perm(__byte_perm(r[4], r[5], 0x0040), __byte_perm(r[6], r[7], 0x0040), 0x5410);
        }
        *reinterpret_cast<uint2 *>(P.dst_alpha + b * P.da_bs + (size_t) line * w + x) = make_uint2(ao[0], ao[1]);
    }
}

// yuv422 -> yuv420p : filter_imageconvert.c:327-368
__global__ void __launch_bounds__(256) k_yuv422_420p(CvtParams P)
{
    const int w = P.width, h = P.height;
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    long long total = (long long) w * h * P.batch;
    if (gid >= total)
        return;
    int x = (int) (gid % w);
    long long t = gid / w;
    int line = (int) (t % h);
    int b = (int) (t / h);
    const uint8_t *src = P.src + b * P.src_bs;
    uint8_t *dst = P.dst + b * P.dst_bs;
    dst[(size_t) line * w + x] = src[line * P.src_stride + 2 * (size_t) x];
    if (line < h / 2 && x < w / 2) {
        uint8_t *pu = dst + (size_t) w * h, *pv = pu + (size_t) (w / 2) * (h / 2);
        pu[(size_t) line * (w / 2) + x] = src[(size_t) line * 2 * P.src_stride + 4 * (size_t) x + 1];
        pv[(size_t) line * (w / 2) + x] = src[(size_t) line * 2 * P.src_stride + 4 * (size_t) x + 3];
    }
}

// The same for 16-pixel groups (w % 16 == 0, 16-byte aligned rows and planes): 32 bytes in, one
// 16-byte Y store and, on the even lines, 8 + 8 chroma bytes; 2-D grid, no div/mod.
__global__ void __launch_bounds__(256) k_yuv422_420p_v(CvtParams P)
{
    const int g = blockIdx.x * 32 + threadIdx.x, line = blockIdx.y * 8 + threadIdx.y;
    const int w = P.width, h = P.height;
    if (g >= (w >> 4) || line >= h)
        return;
    const int b = blockIdx.z;
    const uint8_t *src = P.src + b * P.src_bs + (size_t) line * P.src_stride + 32 * (size_t) g;
    uint8_t *dst = P.dst + b * P.dst_bs;
    const uint4 a = ldg16(src), c = ldg16(src + 16);
    const uint32_t wv[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    uint32_t y[4];
#pragma unroll
    for (int k = 0; k < 4; k++)
        y[k] = __byte_perm(wv[2 * k], wv[2 * k + 1], 0x6420);
    *reinterpret_cast<uint4 *>(dst + (size_t) line * w + 16 * (size_t) g) = make_uint4(y[0], y[1], y[2], y[3]);
    if (!(line & 1) && (line >> 1) < h / 2) { // chroma of the even lines only (:352-366)
        uint32_t t[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            t[k] = __byte_perm(wv[2 * k], wv[2 * k + 1], 0x7351); // u, u', v, v'
        uint8_t *pu = dst + (size_t) w * h, *pv = pu + (size_t) (w / 2) * (h / 2);
        const size_t off = (size_t) (line >> 1) * (w / 2) + 8 * (size_t) g;
        *reinterpret_cast<uint2 *>(pu + off) = make_uint2(__byte_perm(t[0], t[1], 0x5410), __byte_perm(t[2], t[3], 0x5410));
        *reinterpret_cast<uint2 *>(pv + off) = make_uint2(__byte_perm(t[0], t[1], 0x7632), __byte_perm(t[2], t[3], 0x7632));
    }
}

int bytes_pp(int f)
{
    switch (f) {
    case MLTCUDA_FMT_RGB: return 3;
    case MLTCUDA_FMT_RGBA: return 4;
    case MLTCUDA_FMT_YUV422: return 2;
    case MLTCUDA_FMT_RGBA64: return 8;
    case MLTCUDA_FMT_YUV420P: return 1;
    }
    return 0;
}

template <int CELL> int launch_pair(mltcuda_ctx *ctx, const CvtParams &P)
{
    long long total = (long long) ((P.width + 1) / 2) * P.height * P.batch;
    if (total == 0)
        return 0;
    k_convert_pair<CELL><<<mc_blocks(total, 256), 256, 0, ctx->stream>>>(P);
    return mc_after_launch(ctx, "k_convert_pair");
}

} // namespace

extern "C" int mltcuda_convert(mltcuda_ctx *ctx, const mltcuda_convert_desc *d)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!d || !d->src || !d->dst, "null image pointer");
    MC_BADARG(d->width < 0 || d->height < 0, "negative size");
    const int sf = d->src_format, df = d->dst_format;
    const int sb = bytes_pp(sf), db = bytes_pp(df);
    if (!sb || !db || sf == df) {
        mltcuda_set_error("imageconvert: no conversion from format %d to %d", sf, df);
        return MLTCUDA_E_UNSUPPORTED;
    }
    CvtParams P;
    P.src = (const uint8_t *) d->src, P.dst = (uint8_t *) d->dst;
    P.src_alpha = d->src_alpha, P.dst_alpha = d->dst_alpha;
    P.width = d->width, P.height = d->height;
    P.src_stride = d->src_stride ? d->src_stride : (long) d->width * sb;
    P.dst_stride = d->dst_stride ? d->dst_stride : (long) d->width * db;
    P.batch = d->batch > 0 ? d->batch : 1;
    P.src_bs = d->src_bstride, P.dst_bs = d->dst_bstride;
    P.sa_bs = d->src_alpha_bstride, P.da_bs = d->dst_alpha_bstride;
    if (P.batch > 1)
        MC_BADARG(!P.src_bs || !P.dst_bs, "batch > 1 needs src/dst bstride");
    if (d->width == 0 || d->height == 0)
        return 0;
    if ((sf == MLTCUDA_FMT_RGBA || sf == MLTCUDA_FMT_RGBA64)
        && (df == MLTCUDA_FMT_YUV422 || df == MLTCUDA_FMT_RGB))
        MC_BADARG(!d->dst_alpha, "rgba source needs dst_alpha (mlt_image_alloc_alpha in the reference)");

    const bool tight = P.src_stride == (long) d->width * sb && P.dst_stride == (long) d->width * db;
    const bool aligned = (((uintptr_t) P.src | (uintptr_t) P.dst | P.src_bs | P.dst_bs) & 15) == 0;
    // vector tier: even width, 16-byte aligned rows (or a tight frame treated as one long row)
    if (aligned && d->width % 2 == 0
        && ((sf == MLTCUDA_FMT_YUV422 && df == MLTCUDA_FMT_RGBA) || (sf == MLTCUDA_FMT_RGBA && df == MLTCUDA_FMT_YUV422))) {
        bool rows_ok = (P.src_stride % 16 == 0) && (P.dst_stride % 16 == 0);
        if (tight || rows_ok) {
            if (tight)
                P.rows = 1, P.row_px = d->width * d->height;
            else
                P.rows = d->height, P.row_px = d->width;
            // alpha planes are indexed row*row_px: only valid when rows are tight in the plane,
            // which holds for both modes (alpha planes are always width bytes per row).
            long long total = (long long) ((P.row_px + 7) / 8) * P.rows * P.batch;
            if (sf == MLTCUDA_FMT_YUV422)
                k_yuv422_rgba_v<<<mc_blocks(total, 256), 256, 0, ctx->stream>>>(P);
            else
                k_rgba_yuv422_v<<<mc_blocks(total, 256), 256, 0, ctx->stream>>>(P);
            return mc_after_launch(ctx, "k_convert_v");
        }
    }
    // vector tier of the other cells: tight frames, width % 8 (16 from yuv420p), aligned planes
    const bool vec_ok = aligned && tight && d->width % (sf == MLTCUDA_FMT_YUV420P ? 16 : 8) == 0
                        && (((uintptr_t) P.src_alpha | (uintptr_t) P.dst_alpha | P.sa_bs | P.da_bs) & 7) == 0
                        && (d->height + 7) / 8 <= 65535 && P.batch <= 65535
                        && (sf != MLTCUDA_FMT_YUV420P || (((size_t) d->width * d->height) % 4 == 0 && ((size_t) (d->width / 2) * (d->height / 2)) % 4 == 0));
#define CELLCASE(S, D, C) \
    if (sf == MLTCUDA_FMT_##S && df == MLTCUDA_FMT_##D) { \
        if (vec_ok && C != C_YUV422_RGBA && C != C_RGBA_YUV422) { \
            dim3 blk(32, 8, 1), grd((d->width / 8 + 31) / 32, (d->height + 7) / 8, P.batch); \
            k_convert_vec<C><<<grd, blk, 0, ctx->stream>>>(P); \
            return mc_after_launch(ctx, "k_convert_vec"); \
        } \
        return launch_pair<C>(ctx, P); \
    }
    CELLCASE(YUV422, RGB, C_YUV422_RGB)
    CELLCASE(YUV422, RGBA, C_YUV422_RGBA)
    CELLCASE(RGB, YUV422, C_RGB_YUV422)
    CELLCASE(RGBA, YUV422, C_RGBA_YUV422)
    CELLCASE(YUV420P, YUV422, C_420_422)
    CELLCASE(YUV420P, RGB, C_420_RGB)
    CELLCASE(YUV420P, RGBA, C_420_RGBA)
    CELLCASE(RGB, RGBA, C_RGB_RGBA)
    CELLCASE(RGBA, RGB, C_RGBA_RGB)
    CELLCASE(RGBA, RGBA64, C_RGBA_RGBA64)
    CELLCASE(RGBA64, RGBA, C_RGBA64_RGBA)
    CELLCASE(YUV422, RGBA64, C_YUV422_RGBA64)
    CELLCASE(RGBA64, YUV422, C_RGBA64_YUV422)
#undef CELLCASE
    if (sf == MLTCUDA_FMT_YUV422 && df == MLTCUDA_FMT_YUV420P) {
        if (aligned && d->width % 16 == 0 && P.src_stride % 16 == 0 && ((size_t) d->width * d->height) % 16 == 0
            && (((size_t) (d->width / 2) * (d->height / 2)) % 8) == 0 && (d->height + 7) / 8 <= 65535 && P.batch <= 65535) {
            dim3 blk(32, 8, 1), grd((d->width / 16 + 31) / 32, (d->height + 7) / 8, P.batch);
            k_yuv422_420p_v<<<grd, blk, 0, ctx->stream>>>(P);
            return mc_after_launch(ctx, "k_yuv422_420p_v");
        }
        long long total = (long long) d->width * d->height * P.batch;
        k_yuv422_420p<<<mc_blocks(total, 256), 256, 0, ctx->stream>>>(P);
        return mc_after_launch(ctx, "k_yuv422_420p");
    }
    // NULL cells of the reference matrix (:423-437): "imageconvert: no conversion from %s to %s"
    mltcuda_set_error("imageconvert: no conversion from format %d to %d", sf, df);
    return MLTCUDA_E_UNSUPPORTED;
}
