// producers.cu -- device-side synthetic producers and normalisers, so the
// benchmark configurations' sources never cross PCIe:
//   noise  : reference src/modules/core/producer_noise.c:37-51,108-123
//   colour : reference src/modules/core/producer_colour.c:140-205,228-236
//   rescale: reference src/modules/core/filter_rescale.c:48-120 (nearest-neighbour yuv422)
//
// The noise producer's two 16-bit multiply-with-carry generators are sequential in the reference
// (and the buffer is filled from its END backwards).  A lag-1 MWC with multiplier a and base 2^16
// satisfies t[n+1] = a*t[n] mod (a*2^16-1) once the state is below the modulus (true from the 2nd
// step on), so the state at step n is a^(n-2)*t[2] mod m: each thread jumps straight to the state
// of the last pixel of its 64-pixel group with three table look-ups (a^n split into 12+12+8 bit
// digits) and then steps the generator 64 times exactly like the reference.  Bit-exact.
#include "common.cuh"
#include "pixel_ops.cuh"
#include <mutex>

namespace {

constexpr uint32_t MWC_A = 18000u, MWC_B = 30903u;
constexpr uint64_t MOD_A = (uint64_t) MWC_A * 65536u - 1, MOD_B = (uint64_t) MWC_B * 65536u - 1;
// table layout (uint32): [gen][level0: 4096][level1: 4096][level2: 256]
constexpr int TAB_PER_GEN = 4096 + 4096 + 256;

// r mod m for r < 2^63, m < 2^32, with M = floor(2^64 / m): Barrett, quotient off by at most 2
__device__ __forceinline__ uint32_t barrett(uint64_t r, uint64_t m, uint64_t M)
{
    uint64_t t = r - __umul64hi(r, M) * m;
    t = t >= m ? t - m : t;
    t = t >= m ? t - m : t;
    return (uint32_t) t;
}

__device__ __forceinline__ uint32_t mwc_pow(const uint32_t *tab, uint64_t mod, uint64_t M, uint32_t n)
{
    uint64_t r = __ldg(tab + (n & 4095));
    r = barrett(r * __ldg(tab + 4096 + ((n >> 12) & 4095)), mod, M);
    r = barrett(r * __ldg(tab + 8192 + ((n >> 24) & 255)), mod, M);
    return (uint32_t) r;
}

// The pixel byte is ((x << 16) + (y & 65535)) & 0xff == y & 0xff (producer_noise.c:118): the x
// generator never reaches the image, so only the y generator (multiplier 30903) is evaluated.
// One thread = 64 pixels (eight 16-byte stores); grid = (groups of a frame, 1, frame).
constexpr int NOISE_PPT = 64, NOISE_THREADS = 128; // PPT / 8 must be a power of two <= 8
__global__ void __launch_bounds__(NOISE_THREADS) k_noise(uint8_t *dst, unsigned npx, uint32_t y2, uint32_t y0,
                                               const uint32_t *tab, size_t bs, int position, uint64_t M)
{
    const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
    // every lane of the warp owns a whole group: stores go through shared memory (below)
    const bool fullwarp = __ballot_sync(0xffffffffu, ((unsigned long long) c + 1) * NOISE_PPT <= npx) == 0xffffffffu;
    if ((unsigned long long) c * NOISE_PPT >= npx)
        return;
    const int b = blockIdx.z;
    if (b) { // batch frame b is timeline position `position + b`: re-seed (init_seed :37-43)
        uint32_t p = (uint32_t) (position + b);
        y0 = 362436069u - p + (p << 16);
        uint32_t y = y0;
        for (int k = 0; k < 2; k++)
            y = MWC_B * (y & 65535) + (y >> 16);
        y2 = y;
    }
    const unsigned lo = c * NOISE_PPT;
    const unsigned hi = lo + NOISE_PPT - 1 < npx ? lo + NOISE_PPT - 1 : npx - 1; // last pixel of the group
    const unsigned n = npx - hi; // its generation step (1-based; the buffer is filled from the end)
    uint32_t y;
    if (n >= 2)
        y = barrett((uint64_t) mwc_pow(tab + TAB_PER_GEN, MOD_B, M, n - 2) * y2, MOD_B, M);
    else
        y = MWC_B * (y0 & 65535) + (y0 >> 16);
    uint32_t w[NOISE_PPT / 2];
    const int cnt = (int) (hi - lo) + 1;
    uint8_t *d = dst + b * bs + (size_t) lo * 2;
    if (cnt == NOISE_PPT) { // whole group: no per-pixel predicates
#pragma unroll
        for (int k = NOISE_PPT / 2 - 1; k >= 0; k--) {
            // clamp(v, 16, 240) - 16 = relu(min(v - 16, 224)), one VIADDMNMX.RELU; the + 16 and
            // U/V = 128 are added per pair
            int v0, v1;
            asm("min.relu.s32 %0, %1, 224;" : "=r"(v1) : "r"((int) (y & 0xff) - 16));
            y = MWC_B * (y & 65535) + (y >> 16);
            asm("min.relu.s32 %0, %1, 224;" : "=r"(v0) : "r"((int) (y & 0xff) - 16));
            y = MWC_B * (y & 65535) + (y >> 16);
            w[k] = __byte_perm(v0, v1, 0x1410) + 0x80108010u; // byte 1 of v0 is zero
        }
        constexpr int U = NOISE_PPT / 8; // 16-byte units per thread
        if (fullwarp) {
            // A thread's 128 bytes are contiguous, so a direct 16-byte store instruction would touch
            // 32 separate 128-byte-strided pieces.  Transpose through shared memory (16-byte units,
            // XOR-swizzled: conflict-free both ways) so each store instruction writes 512
            // contiguous bytes of the warp's 4 KiB.
            __shared__ uint4 s_t[NOISE_THREADS / 32][32 * U];
            uint4 *sm = s_t[threadIdx.x >> 5];
            const int l = threadIdx.x & 31;
#pragma unroll
            for (int k = 0; k < U; k++)
                sm[U * l + (k ^ (l & (U - 1)))] = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
            __syncwarp();
            uint4 *dw = reinterpret_cast<uint4 *>(d - (size_t) l * (NOISE_PPT * 2)); // the warp's first byte
#pragma unroll
            for (int k = 0; k < U; k++) {
                const int i = 32 * k + l, lp = i / U;
                dw[i] = sm[U * lp + ((i & (U - 1)) ^ (lp & (U - 1)))];
            }
            return;
        }
#pragma unroll
        for (int k = 0; k < U; k++)
            reinterpret_cast<uint4 *>(d)[k] = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
        return;
    }
#pragma unroll
    for (int k = 0; k < NOISE_PPT / 2; k++)
        w[k] = 0;
#pragma unroll
    for (int k = NOISE_PPT - 1; k >= 0; k--) {
        if (k < cnt) {
            uint32_t value = y & 0xff;
            value = value < 16 ? 16 : value > 240 ? 240 : value;
            w[k >> 1] |= (value | (128u << 8)) << ((k & 1) * 16);
            y = MWC_B * (y & 65535) + (y >> 16);
        }
    }
#pragma unroll
    for (int k = 0; k < NOISE_PPT; k++)
        if (k < cnt)
            reinterpret_cast<uint16_t *>(d)[k] = (uint16_t) (w[k >> 1] >> ((k & 1) * 16));
}

__global__ void __launch_bounds__(256) k_fill16(uint4 *dst, size_t n16, uint4 v)
{
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n16)
        dst[i] = v;
}

__global__ void __launch_bounds__(256) k_colour_yuv422(uint8_t *dst, int width, int height, uint32_t y, uint32_t u, uint32_t v)
{
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    int ppr = (width + 1) / 2;
    if (gid >= (long long) ppr * height)
        return;
    int pi = (int) (gid % ppr), line = (int) (gid / ppr);
    uint8_t *d = dst + (size_t) line * width * 2 + 4 * (size_t) pi;
    d[0] = y, d[1] = u;
    if (2 * pi + 1 < width)
        d[2] = y, d[3] = v;
}

// filter_scale (:48-120)
// One thread = 4 output pixel pairs (16 bytes) of one row; grid = (quads of a row, row, frame).
__global__ void __launch_bounds__(256) k_rescale_yuv422(uint8_t *dst, int ow, const uint8_t *src, int istride,
                                                        long in_middle, int scale_width, int scale_height, int outer,
                                                        int bottom, int ncols, size_t dst_bs, size_t src_bs)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q * 4 >= ncols)
        return;
    const int row = blockIdx.y;
    const int dy = -bottom + row * scale_height;
    const uint8_t *in_line = src + blockIdx.z * src_bs + in_middle + (long) (dy >> 16) * istride;
    uint8_t *o = dst + blockIdx.z * dst_bs + (size_t) row * ow * 2 + 16 * (size_t) q;
    uint32_t r[4] = {0, 0, 0, 0};
    const int npair = ncols - q * 4 < 4 ? ncols - q * 4 : 4;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (k < npair) {
            int dx = -outer + 2 * (q * 4 + k) * scale_width; // :97-109, two pixels per trip
            int base = dx >> 15;
            base &= 0xfffffffe;
            uint32_t v = __ldg(in_line + base);
            base &= 0xfffffffc;
            v |= (uint32_t) __ldg(in_line + base + 1) << 8;
            dx += scale_width;
            base = dx >> 15;
            base &= 0xfffffffe;
            v |= (uint32_t) __ldg(in_line + base) << 16;
            base &= 0xfffffffc;
            v |= (uint32_t) __ldg(in_line + base + 3) << 24;
            r[k] = v;
        }
    if (npair == 4 && ((uintptr_t) o & 15) == 0) {
        *reinterpret_cast<uint4 *>(o) = make_uint4(r[0], r[1], r[2], r[3]);
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (k < npair)
                reinterpret_cast<uint32_t *>(o)[k] = r[k]; // rows start on 4-byte boundaries (ow even)
    }
}

// scale_alpha (filter_rescale.c:122-147): nearest neighbour sampled at the pixel centres
__global__ void __launch_bounds__(256) k_rescale_alpha(uint8_t *dst, int ow, int oh, const uint8_t *src, int iw, int ox, int oy)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= ow)
        return;
    const int y = (oy >> 1) + i * oy, x = (ox >> 1) + j * ox;
    dst[(size_t) i * ow + j] = __ldg(src + (size_t) (y >> 16) * iw + (x >> 16));
}

// resize_image / resize_alpha (filter_resize.c:30-137): the input centred in a frame filled with the
// format's "nothing" (pattern = the fill bytes, period 8).  One thread = 4 output bytes.
__global__ void __launch_bounds__(256) k_resize_pad(uint8_t *dst, int orow, int oh, const uint8_t *src, int irow, int ih,
                                                    int off_x, int off_y, unsigned long long pattern)
{
    const int xb = (blockIdx.x * blockDim.x + threadIdx.x) * 4, row = blockIdx.y;
    if (xb >= orow)
        return;
    const int ir = row - off_y;
    const uint8_t *in = src + (long) ir * irow - off_x;
    uint8_t *o = dst + (size_t) row * orow + xb;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int x = xb + k;
        uint32_t byte = (uint32_t) (pattern >> (8 * (x & 7))) & 0xff;
        if (ir >= 0 && ir < ih && x >= off_x && x < off_x + irow)
            byte = __ldg(in + x);
        v |= byte << (8 * k);
    }
    if (xb + 4 <= orow && ((uintptr_t) o & 3) == 0) {
        *reinterpret_cast<uint32_t *>(o) = v;
    } else {
        for (int k = 0; k < 4 && xb + k < orow; k++)
            o[k] = (uint8_t) (v >> (8 * k));
    }
}

// mirror:flip (src/modules/core/filter_mirror.c:247-281): every row reversed pair-wise; inside a pair the two
// lumas swap and the shared chroma stays (output pair k = input pair n-1-k as Y1 U Y0 V).  One thread = VEC pairs.
template <int VEC> __global__ void __launch_bounds__(256) k_mirror_flip_yuv422(uint8_t *dst, const uint8_t *src, int npairs)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x; // group of VEC output pairs
    if (g * VEC >= npairs)
        return;
    const size_t row = (size_t) blockIdx.y * npairs * 4;
    const uint32_t *in = reinterpret_cast<const uint32_t *>(src + row);
    uint32_t *out = reinterpret_cast<uint32_t *>(dst + row);
    if (VEC == 4) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(in + (npairs - 4 - 4 * g)));
        *reinterpret_cast<uint4 *>(out + 4 * g) = make_uint4(__byte_perm(v.w, 0, 0x3012), __byte_perm(v.z, 0, 0x3012),
                                                             __byte_perm(v.y, 0, 0x3012), __byte_perm(v.x, 0, 0x3012));
    } else {
        out[g] = __byte_perm(__ldg(in + (npairs - 1 - g)), 0, 0x3012);
    }
}
// the alpha plane of the same filter (:283-294): every row reversed
__global__ void __launch_bounds__(256) k_mirror_flip_plane(uint8_t *dst, const uint8_t *src, int w)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < w)
        dst[(size_t) blockIdx.y * w + j] = __ldg(src + (size_t) blockIdx.y * w + (w - 1 - j));
}

int ensure_noise_tab(mltcuda_ctx *ctx)
{
    if (ctx->noise_tab)
        return 0;
    static uint32_t host[2 * TAB_PER_GEN];
    const uint64_t mods[2] = {MOD_A, MOD_B};
    const uint64_t as[2] = {MWC_A, MWC_B};
    for (int g = 0; g < 2; g++) {
        uint32_t *t = host + g * TAB_PER_GEN;
        uint64_t m = mods[g], v = 1;
        for (int i = 0; i < 4096; i++)
            t[i] = (uint32_t) v, v = v * as[g] % m; // v = a^4096 at the end
        uint64_t s1 = v;
        v = 1;
        for (int i = 0; i < 4096; i++)
            t[4096 + i] = (uint32_t) v, v = v * s1 % m; // v = a^(2^24)
        uint64_t s2 = v;
        v = 1;
        for (int i = 0; i < 256; i++)
            t[8192 + i] = (uint32_t) v, v = v * s2 % m;
    }
    MC_CHECK(cudaMalloc(&ctx->noise_tab, sizeof(host)));
    MC_CHECK(cudaMemcpy(ctx->noise_tab, host, sizeof(host), cudaMemcpyHostToDevice));
    return 0;
}

} // namespace

extern "C" int mltcuda_noise(mltcuda_ctx *ctx, uint8_t *dst, int width, int height, int position, int batch,
                             size_t dst_bstride)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || width < 0 || height < 0, "bad argument");
    MC_BADARG(((uintptr_t) dst & 15) || (dst_bstride & 15), "noise destination must be 16-byte aligned");
    long long npx = (long long) width * height;
    if (npx == 0)
        return 0;
    if (batch <= 0)
        batch = 1;
    if (int e = ensure_noise_tab(ctx))
        return e;
    uint32_t p = (uint32_t) position;
    uint32_t x0 = 521288629u + p - (p << 16), y0 = 362436069u - p + (p << 16);
    uint32_t x = x0, y = y0;
    for (int k = 0; k < 2; k++)
        x = MWC_A * (x & 65535) + (x >> 16), y = MWC_B * (y & 65535) + (y >> 16);
    MC_BADARG(npx >= (1ll << 31) || batch > 65535, "frame too large / too many frames for one launch");
    (void) x, (void) x0;
    dim3 grd(mc_blocks((npx + NOISE_PPT - 1) / NOISE_PPT, NOISE_THREADS), 1, batch);
    k_noise<<<grd, NOISE_THREADS, 0, ctx->stream>>>(dst, (unsigned) npx, y, y0, ctx->noise_tab, dst_bstride, position, ~0ull / MOD_B);
    return mc_after_launch(ctx, "k_noise");
}

extern "C" int mltcuda_colour(mltcuda_ctx *ctx, void *dst, uint8_t *dst_alpha, int format, int width, int height,
                              uint8_t r, uint8_t g, uint8_t b, uint8_t a)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || width < 0 || height < 0, "bad argument");
    size_t n = (size_t) width * height;
    if (n == 0)
        return 0;
    if (format == MLTCUDA_FMT_YUV422) {
        int y, u, v;
        y = ((263 * r + 516 * g + 100 * b) >> 10) + 16;
        u = ((-152 * r - 300 * g + 450 * b) >> 10) + 128;
        v = ((450 * r - 377 * g - 73 * b) >> 10) + 128;
        long long total = (long long) ((width + 1) / 2) * height;
        k_colour_yuv422<<<mc_blocks(total, 256), 256, 0, ctx->stream>>>((uint8_t *) dst, width, height, y & 0xff, u & 0xff, v & 0xff);
        if (int e = mc_after_launch(ctx, "k_colour_yuv422"))
            return e;
        if (dst_alpha)
            MC_CHECK(cudaMemsetAsync(dst_alpha, a, n, ctx->stream));
        return 0;
    }
    uint4 v;
    size_t bytes;
    if (format == MLTCUDA_FMT_RGBA) {
        uint32_t p = r | (g << 8) | (b << 16) | ((uint32_t) a << 24);
        v = make_uint4(p, p, p, p), bytes = n * 4;
    } else if (format == MLTCUDA_FMT_RGBA64) {
        uint32_t lo = ((r << 8) + r) | (((g << 8) + g) << 16), hi = ((b << 8) + b) | ((uint32_t) ((a << 8) + a) << 16);
        v = make_uint4(lo, hi, lo, hi), bytes = n * 8;
    } else {
        mltcuda_set_error("colour: unsupported format %d", format);
        return MLTCUDA_E_UNSUPPORTED;
    }
    MC_BADARG((uintptr_t) dst & 15, "colour destination must be 16-byte aligned");
    size_t n16 = (bytes + 15) / 16; // tail slack of mltcuda_malloc absorbs the round-up
    k_fill16<<<mc_blocks(n16, 256), 256, 0, ctx->stream>>>((uint4 *) dst, n16, v);
    return mc_after_launch(ctx, "k_fill16");
}

extern "C" int mltcuda_rescale_yuv422(mltcuda_ctx *ctx, uint8_t *dst, int ow, int oh, const uint8_t *src, int iw, int ih,
                                      int batch, size_t dst_bstride, size_t src_bstride)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || !src || ow <= 0 || oh <= 0 || iw <= 0 || ih <= 0, "bad argument");
    MC_BADARG((ow & 1) || ((uintptr_t) dst & 3) || (dst_bstride & 3), "rescale: even output width, 4-byte aligned destination");
    if (batch <= 0)
        batch = 1;
    MC_BADARG(batch > 1 && (!dst_bstride || !src_bstride), "batch > 1 needs both bstrides");
    int istride = iw * 2;
    iw -= iw % 4;
    int out_x_range = ow / 2, out_y_range = oh / 2, in_x_range = iw / 2, in_y_range = ih / 2;
    long in_middle = (long) istride * in_y_range + in_x_range * 2;
    int scale_width = (iw << 16) / ow, scale_height = (ih << 16) / oh;
    int outer = out_x_range * scale_width, bottom = out_y_range * scale_height;
    if (scale_width <= 0 || scale_height <= 0 || outer <= 0 || bottom <= 0)
        return 0;
    // loop trip counts of :93-110: dy from -bottom step scale_height while < bottom; dx advances twice per pair
    int nrows = (2 * bottom + scale_height - 1) / scale_height;
    int ncols = (2 * outer + 2 * scale_width - 1) / (2 * scale_width);
    MC_BADARG(nrows > 65535 || batch > 65535, "more than 65535 rows / frames in one launch");
    dim3 grd(mc_blocks((ncols + 3) / 4, 128), nrows, batch);
    k_rescale_yuv422<<<grd, 128, 0, ctx->stream>>>(dst, ow, src, istride, in_middle, scale_width, scale_height, outer, bottom,
                                                   ncols, dst_bstride, src_bstride);
    return mc_after_launch(ctx, "k_rescale_yuv422");
}

extern "C" int mltcuda_rescale_alpha(mltcuda_ctx *ctx, uint8_t *dst, int ow, int oh, const uint8_t *src, int iw, int ih)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || !src || ow <= 0 || oh <= 0 || iw <= 0 || ih <= 0 || oh > 65535, "bad argument");
    dim3 grd(mc_blocks(ow, 256), oh, 1);
    k_rescale_alpha<<<grd, 256, 0, ctx->stream>>>(dst, ow, oh, src, iw, (iw << 16) / ow, (ih << 16) / oh);
    return mc_after_launch(ctx, "k_rescale_alpha");
}

extern "C" int mltcuda_resize_pad(mltcuda_ctx *ctx, void *dst, int ow, int oh, const void *src, int iw, int ih, int format,
                                  int alpha_value)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || !src || ow <= 0 || oh <= 0 || iw <= 0 || ih <= 0 || iw > ow || ih > oh || oh > 65535, "bad argument");
    int bpp;
    unsigned long long pattern = 0;
    const unsigned long long a = (unsigned) alpha_value & 0xff;
    switch (format) {
    case MLTCUDA_FMT_NONE: // a bare alpha plane (resize_alpha :30-63)
        bpp = 1, pattern = a * 0x0101010101010101ull;
        break;
    case MLTCUDA_FMT_YUV422: // Y = 16, U = V = 128 (:108-114)
        bpp = 2, pattern = 0x8010801080108010ull;
        break;
    case MLTCUDA_FMT_RGBA: // transparent black, alpha = resize_alpha (:90-97)
        bpp = 4, pattern = (a << 24) | (a << 56);
        break;
    case MLTCUDA_FMT_RGBA64: // 16-bit alpha = alpha << 8, little endian (:98-107)
        bpp = 8, pattern = a << 56;
        break;
    case MLTCUDA_FMT_RGB:
        bpp = 3;
        break;
    default:
        mltcuda_set_error("resize_pad: unsupported format %d", format);
        return MLTCUDA_E_UNSUPPORTED;
    }
    int off_x = (ow - iw) / 2 * bpp, off_y = (oh - ih) / 2;
    if (format == MLTCUDA_FMT_YUV422)
        off_x -= off_x % 4; // keep the chroma phase (:114)
    if (format == MLTCUDA_FMT_NONE)
        off_x -= off_x % 2; // (:42)
    dim3 grd(mc_blocks(((long long) ow * bpp + 3) / 4, 256), oh, 1);
    k_resize_pad<<<grd, 256, 0, ctx->stream>>>((uint8_t *) dst, ow * bpp, oh, (const uint8_t *) src, iw * bpp, ih, off_x, off_y,
                                               pattern);
    return mc_after_launch(ctx, "k_resize_pad");
}

extern "C" int mltcuda_mirror_flip(mltcuda_ctx *ctx, uint8_t *dst, uint8_t *dst_alpha, const uint8_t *src, const uint8_t *src_alpha,
                                   int width, int height)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || !src || dst == src || width <= 0 || height <= 0 || height > 65535, "bad argument");
    if (width % 2) {
        mltcuda_set_error("mirror flip: odd width %d (the reference's odd-width byte shuffle is not reproduced on the device)", width);
        return MLTCUDA_E_UNSUPPORTED;
    }
    const int npairs = width / 2;
    if (npairs % 4 == 0 && (((uintptr_t) dst | (uintptr_t) src) & 15) == 0)
        k_mirror_flip_yuv422<4><<<dim3(mc_blocks(npairs / 4, 256), height, 1), 256, 0, ctx->stream>>>(dst, src, npairs);
    else
        k_mirror_flip_yuv422<1><<<dim3(mc_blocks(npairs, 256), height, 1), 256, 0, ctx->stream>>>(dst, src, npairs);
    if (int e = mc_after_launch(ctx, "k_mirror_flip_yuv422"))
        return e;
    if (src_alpha && dst_alpha) {
        k_mirror_flip_plane<<<dim3(mc_blocks(width, 256), height, 1), 256, 0, ctx->stream>>>(dst_alpha, src_alpha, width);
        return mc_after_launch(ctx, "k_mirror_flip_plane");
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// mltcuda_noise_static: the noise producer for callers that must not allocate (include/mltcuda.h).
//
// mltcuda_noise keeps its skip-ahead tables in a cudaMalloc'ed block per context.  A render thread in
// the middle of real-time playback cannot afford that: cudaMalloc / cudaFree take a device-wide lock
// and wait for the GPU to drain, which stalls every other worker's stream for a frame or more; and a
// host application that has handed all device memory to its own pool (or runs inside a fixed MIG / MPS
// memory budget) may get cudaErrorMemoryAllocation at a moment it cannot handle.  Here the tables sit
// in a file-scope __device__ array instead: the module image carries them, every GPU that loads this
// library holds MLTCUDA_NOISE_STATIC_BYTES of it until the process exits, and no call -- the first
// one included -- asks CUDA or any library for memory.
//
// On the host, `g_noise_tab_static` is only a handle for the symbol, not a device address; the kernel
// must be given what cudaGetSymbolAddress returns for the CURRENT device (each GPU has its own copy).
// That lookup and the one-time fill of the tables happen in the first call on each device, under a
// mutex, with a blocking copy on the caller's stream -- so make that first call outside any stream
// capture and before latency matters.  Afterwards the array is read-only: calls on any number of
// contexts of the same GPU may overlap on the GPU without racing (they only read it), nothing has to
// be zeroed between calls, and what one call leaves in the array is exactly what the next one needs.
// The fill itself cannot race with a kernel either: the address is handed out only after the copy
// has completed.  (A variant whose kernels WROTE to shared static scratch would have to serialise.)
namespace {

constexpr int NOISE_STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) uint32_t g_noise_tab_static[2 * TAB_PER_GEN];
static_assert(sizeof(g_noise_tab_static) == MLTCUDA_NOISE_STATIC_BYTES, "header and table size disagree");

std::mutex g_noise_static_mu;
const uint32_t *g_noise_static_ptr[NOISE_STATIC_MAX_DEVICES]; // device address of the array, per GPU; null = not set up

// powers of a modulo m in three digit levels (12 + 12 + 8 bits of the exponent), as k_noise's mwc_pow expects
void noise_static_fill(uint32_t *t, uint64_t a, uint64_t m)
{
    const int level_len[3] = {4096, 4096, 256};
    uint64_t base = a;
    for (int level = 0; level < 3; level++) {
        uint64_t v = 1;
        for (int i = 0; i < level_len[level]; i++)
            *t++ = (uint32_t) v, v = v * base % m;
        base = v; // base^len: the next level's unit
    }
}

// Device address of the static tables on ctx's GPU (which mc_use has made current); fills them on first use.
int noise_static_table(mltcuda_ctx *ctx, const uint32_t **out)
{
    if (ctx->device < 0 || ctx->device >= NOISE_STATIC_MAX_DEVICES) {
        mltcuda_set_error("mltcuda_noise_static: device %d beyond the %d this entry point tracks", ctx->device, NOISE_STATIC_MAX_DEVICES);
        return MLTCUDA_E_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_noise_static_mu);
    if (!g_noise_static_ptr[ctx->device]) {
        uint32_t *dev = nullptr;
        MC_CHECK(cudaGetSymbolAddress((void **) &dev, g_noise_tab_static));
        static uint32_t host[2 * TAB_PER_GEN]; // guarded by the mutex; same contents for every GPU
        noise_static_fill(host, MWC_A, MOD_A);
        noise_static_fill(host + TAB_PER_GEN, MWC_B, MOD_B);
        MC_CHECK(cudaMemcpyAsync(dev, host, sizeof(host), cudaMemcpyHostToDevice, ctx->stream));
        MC_CHECK(cudaStreamSynchronize(ctx->stream)); // complete before any other context's stream can see the pointer
        g_noise_static_ptr[ctx->device] = dev;
    }
    *out = g_noise_static_ptr[ctx->device];
    return 0;
}

} // namespace

extern "C" int mltcuda_noise_static(mltcuda_ctx *ctx, uint8_t *dst, int width, int height, int position, int batch,
                                    size_t dst_bstride)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || width < 0 || height < 0, "bad argument");
    MC_BADARG(((uintptr_t) dst & 15) || (dst_bstride & 15), "noise destination must be 16-byte aligned");
    const long long npx = (long long) width * height;
    if (npx == 0)
        return 0;
    if (batch <= 0)
        batch = 1;
    if (npx > MLTCUDA_NOISE_STATIC_MAX_PIXELS || batch > MLTCUDA_NOISE_STATIC_MAX_BATCH) {
        mltcuda_set_error("mltcuda_noise_static: %lld pixels x %d frames is beyond what the static tables reach "
                          "(at most %lld pixels per frame, %d frames per call)",
                          npx, batch, (long long) MLTCUDA_NOISE_STATIC_MAX_PIXELS, MLTCUDA_NOISE_STATIC_MAX_BATCH);
        return MLTCUDA_E_BADARG;
    }
    MC_BADARG(batch > 1 && dst_bstride < (size_t) npx * 2, "batch > 1 needs dst_bstride >= width*height*2");
    const uint32_t *tab = nullptr;
    if (int e = noise_static_table(ctx, &tab))
        return e;
    // seed of the y generator for this timeline position and its state after two steps (producer_noise.c:37-51)
    const uint32_t p = (uint32_t) position;
    const uint32_t y0 = 362436069u - p + (p << 16);
    uint32_t y2 = y0;
    for (int k = 0; k < 2; k++)
        y2 = MWC_B * (y2 & 65535) + (y2 >> 16);
    dim3 grd(mc_blocks((npx + NOISE_PPT - 1) / NOISE_PPT, NOISE_THREADS), 1, batch);
    k_noise<<<grd, NOISE_THREADS, 0, ctx->stream>>>(dst, (unsigned) npx, y2, y0, tab, dst_bstride, position, ~0ull / MOD_B);
    return mc_after_launch(ctx, "k_noise");
}
