// affine.cu -- K4: affine transform + sampling + alpha-over (reference
// src/modules/plus/transition_affine.c:342-409 sliced_proc, src/modules/plus/interp.h:50-334).
//
// One thread = 16 bytes of the A image (4 rgba or 2 rgba64 destination pixels).  The inverse
// mapping runs in double exactly as MapX/MapY (:213-227) do, the sampler in float in the
// reference's operation order (-fmad=false).  A is only read (and written back) when at least one
// of the thread's pixels maps inside B, so untouched regions of A cost no HBM traffic; B taps go
// through the read-only path (L1/L2 catch the 4x / 16x tap re-use of neighbouring pixels).
// HBM-bound: 12 B per covered pixel (rgba), 24 B (rgba64).
#include "common.cuh"
#include "pixel_ops.cuh"

namespace {

struct AffK
{
    AffineParams p;
    uint8_t *a_image;
    const uint8_t *b_image;
    int a_width, a_height;
    size_t a_bs, b_bs;
    int batch;
};

template <typename T> __global__ void __launch_bounds__(256) k_affine(AffK K)
{
    constexpr int PPT = sizeof(T) == 1 ? 4 : 2;
    const long long npx = (long long) K.a_width * K.a_height;
    const long long cpf = (npx + PPT - 1) / PPT;
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= cpf * K.batch)
        return;
    int b = (int) (gid / cpf);
    long long p0 = (gid % cpf) * PPT;
    int i = (int) (p0 / K.a_width), j = (int) (p0 % K.a_width);
    const AffineParams &P = K.p;

    // inverse map all pixels first; skip the A traffic when none lands inside B
    double dxs[PPT], dys[PPT];
    int hit = 0;
    {
        int ii = i, jj = j;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            mc_affine_map(P, mc_affine_row(P, ii), jj, dxs[k], dys[k]);
            bool in = p0 + k < npx && mc_affine_inside(P, dxs[k], dys[k]);
            hit |= in ? (1 << k) : 0;
            if (++jj == K.a_width)
                jj = 0, ii++;
        }
    }
    if (!hit)
        return;
    T *q = reinterpret_cast<T *>(K.a_image + b * K.a_bs) + p0 * 4;
    const T *bimg = reinterpret_cast<const T *>(K.b_image + b * K.b_bs);
    const bool full = p0 + PPT <= npx;
    uint32_t qv[PPT][4];
    if (full) {
        uint4 a = ld16(q);
        uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int k = 0; k < PPT; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                if (sizeof(T) == 1)
                    qv[k][c] = (aw[k] >> (8 * c)) & 0xff;
                else
                    qv[k][c] = (aw[2 * k + (c >> 1)] >> (16 * (c & 1))) & 0xffff;
            }
    } else {
#pragma unroll
        for (int k = 0; k < PPT; k++)
#pragma unroll
            for (int c = 0; c < 4; c++)
                qv[k][c] = p0 + k < npx ? q[4 * k + c] : 0;
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        if (hit & (1 << k)) {
            float fx = (float) dxs[k], fy = (float) dys[k];
            if (P.interp == MLTCUDA_INTERP_BL)
                mc_interp_bl<T>(bimg, P.b_width, P.b_height, fx, fy, P.mix, qv[k], P.b_alpha);
            else if (P.interp == MLTCUDA_INTERP_NN)
                mc_interp_nn<T>(bimg, P.b_width, P.b_height, fx, fy, P.mix, qv[k], P.b_alpha);
            else
                mc_interp_bc<T>(bimg, P.b_width, P.b_height, fx, fy, P.mix, qv[k], P.b_alpha);
        }
    }
    if (full) {
        uint32_t ow[4];
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            if (sizeof(T) == 1)
                ow[k] = qv[k][0] | (qv[k][1] << 8) | (qv[k][2] << 16) | (qv[k][3] << 24);
            else
                ow[2 * k] = qv[k][0] | (qv[k][1] << 16), ow[2 * k + 1] = qv[k][2] | (qv[k][3] << 16);
        }
        *reinterpret_cast<uint4 *>(q) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    } else {
#pragma unroll
        for (int k = 0; k < PPT; k++)
#pragma unroll
            for (int c = 0; c < 4; c++)
                if (p0 + k < npx)
                    q[4 * k + c] = (T) qv[k][c];
    }
}


// Row-aligned fast path (a_width % PPT == 0, the usual case): 2-D blocks of 32 groups x 8 rows so a
// CTA's B taps share L1 lines across rows; whole groups are rejected from their two end pixels
// (mc_affine_reject) before any A traffic; INTERP is a template parameter so only one sampler is
// compiled into each instantiation.
template <typename T, int INTERP> __global__ void __launch_bounds__(128) k_affine_rows(AffK K)
{
    constexpr int PPT = sizeof(T) == 1 ? 4 : 2;
    const int g = blockIdx.x * 32 + threadIdx.x;
    const int i = blockIdx.y * 4 + threadIdx.y;
    const int j = g * PPT;
    if (j >= K.a_width || i >= K.a_height)
        return;
    const int b = blockIdx.z;
    const AffineParams &P = K.p;
    if (mc_affine_bb_reject(P, i, j, PPT))
        return;
    const AffineRow R = mc_affine_row(P, i);
    double dxs[PPT], dys[PPT];
    // x of pixel j + k = (lower_x + j) + k: the reference steps x by ++ from lower_x (:361-363), so both
    // are exact for the integer / half-integer origins it produces
    const double x0 = P.lower_x + (double) j;
    mc_affine_map_x(P, R, x0, dxs[0], dys[0]);
    mc_affine_map_x(P, R, x0 + (double) (PPT - 1), dxs[PPT - 1], dys[PPT - 1]);
    if (mc_affine_reject2(P, dxs[0], dys[0], dxs[PPT - 1], dys[PPT - 1]))
        return;
#pragma unroll
    for (int k = 1; k < PPT - 1; k++)
        mc_affine_map_x(P, R, x0 + (double) k, dxs[k], dys[k]);
    // dx(j), dy(j) are monotone along a row, so a group whose two end pixels are inside B is inside
    // as a whole: the interior of the rectangle (nearly all covered groups) skips the per-pixel tests
    // and runs the PPT samplers without branches between them.
    const bool all = mc_affine_inside(P, dxs[0], dys[0]) && mc_affine_inside(P, dxs[PPT - 1], dys[PPT - 1]);
    T *q = reinterpret_cast<T *>(K.a_image + b * K.a_bs) + ((size_t) i * K.a_width + j) * 4;
    const T *bimg = reinterpret_cast<const T *>(K.b_image + b * K.b_bs);
    asm volatile("" : "+l"(bimg)); // keep the batch base in registers: do not re-derive it per pixel
    uint32_t aw[4];
    if (all) {
        uint4 a = ld16(q);
        aw[0] = a.x, aw[1] = a.y, aw[2] = a.z, aw[3] = a.w;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            if constexpr (sizeof(T) == 1) {
                aw[k] = mc_affine_sample_packed<T, INTERP>(P, bimg, dxs[k], dys[k], aw[k]);
            } else {
                uint2 o = mc_affine_sample_packed<T, INTERP>(P, bimg, dxs[k], dys[k], make_uint2(aw[2 * k], aw[2 * k + 1]));
                aw[2 * k] = o.x, aw[2 * k + 1] = o.y;
            }
        }
    } else {
        int hit = 0;
#pragma unroll
        for (int k = 0; k < PPT; k++)
            hit |= mc_affine_inside(P, dxs[k], dys[k]) ? (1 << k) : 0;
        if (!hit)
            return;
        uint4 a = ld16(q);
        aw[0] = a.x, aw[1] = a.y, aw[2] = a.z, aw[3] = a.w;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            if (hit & (1 << k)) {
                if constexpr (sizeof(T) == 1) {
                    aw[k] = mc_affine_sample_packed<T, INTERP>(P, bimg, dxs[k], dys[k], aw[k]);
                } else {
                    uint2 o = mc_affine_sample_packed<T, INTERP>(P, bimg, dxs[k], dys[k], make_uint2(aw[2 * k], aw[2 * k + 1]));
                    aw[2 * k] = o.x, aw[2 * k + 1] = o.y;
                }
            }
        }
    }
    *reinterpret_cast<uint4 *>(q) = make_uint4(aw[0], aw[1], aw[2], aw[3]);
}

template <typename T> static void launch_affine_rows(const AffK &K, int interp, cudaStream_t st)
{
    constexpr int PPT = sizeof(T) == 1 ? 4 : 2;
    dim3 blk(32, 4, 1), grd((K.a_width / PPT + 31) / 32, (K.a_height + 3) / 4, K.batch);
    if (interp == MLTCUDA_INTERP_BL)
        k_affine_rows<T, 1><<<grd, blk, 0, st>>>(K);
    else if (interp == MLTCUDA_INTERP_NN)
        k_affine_rows<T, 0><<<grd, blk, 0, st>>>(K);
    else
        k_affine_rows<T, 2><<<grd, blk, 0, st>>>(K);
}

} // namespace

int mc_fill_affine_params(AffineParams &p, const mltcuda_affine_desc *d)
{
    p.m00 = d->matrix[0][0], p.m01 = d->matrix[0][1], p.m02 = d->matrix[0][2];
    p.m10 = d->matrix[1][0], p.m11 = d->matrix[1][1], p.m12 = d->matrix[1][2];
    p.lower_x = d->lower_x, p.lower_y = d->lower_y, p.dz = d->dz, p.rdz = 1.0 / d->dz;
    p.x_offset = d->x_offset, p.y_offset = d->y_offset;
    p.minima = d->minima, p.xmax = d->xmax, p.ymax = d->ymax;
    p.mix = (float) d->mix;
    p.b_width = d->b_width, p.b_height = d->b_height, p.b_alpha = d->b_alpha, p.interp = d->interp;
    // The reference samplers read a 2x2 (BL) / 4x4 (BC) footprint clamped to the image; smaller
    // images make it read out of bounds.  Refuse instead of faulting.
    // tap offsets are 32-bit unsigned byte offsets (mc_tap2): the reference caps B at 16000 x 16000
    // (MLT_AFFINE_MAX_DIMENSION, transition_affine.c:32), 2 GiB in rgba64
    if ((unsigned long long) d->b_width * (unsigned long long) d->b_height * (d->format == MLTCUDA_FMT_RGBA64 ? 8ull : 4ull) >= (1ull << 32)) {
        mltcuda_set_error("affine: B image %dx%d is 4 GiB or more", d->b_width, d->b_height);
        return MLTCUDA_E_BADARG;
    }
    // ... except bilinear on a 1-wide / 1-high B, which the reference does produce (App. B-3) and
    // whose result is well defined (mc_interp_bl): served by the generic kernel with clamped taps.
    p.degenerate = d->interp == MLTCUDA_INTERP_BL && d->b_width >= 1 && d->b_height >= 1 && (d->b_width < 2 || d->b_height < 2);
    // bounding box of the accepted destination pixels: the four corners of [minima, xmax] x [minima, ymax]
    // mapped back through dx = (m00 x + m01 y + m02) rdz + x_offset, dy likewise
    p.bb_j0 = p.bb_i0 = INT_MIN / 2, p.bb_j1 = p.bb_i1 = INT_MAX / 2;
    const double det = p.m00 * p.m11 - p.m01 * p.m10;
    if (std::isfinite(det) && std::fabs(det) > 1e-12 && std::isfinite(p.rdz) && p.rdz != 0.0) {
        double lo_x = INFINITY, hi_x = -INFINITY, lo_y = INFINITY, hi_y = -INFINITY;
        for (int c = 0; c < 4; c++) {
            const double u = (((c & 1) ? p.xmax : p.minima) - p.x_offset) / p.rdz - p.m02;
            const double v = (((c & 2) ? p.ymax : p.minima) - p.y_offset) / p.rdz - p.m12;
            const double x = (p.m11 * u - p.m01 * v) / det, y = (p.m00 * v - p.m10 * u) / det;
            lo_x = std::fmin(lo_x, x), hi_x = std::fmax(hi_x, x), lo_y = std::fmin(lo_y, y), hi_y = std::fmax(hi_y, y);
        }
        lo_x -= p.lower_x, hi_x -= p.lower_x, lo_y -= p.lower_y, hi_y -= p.lower_y;
        if (std::isfinite(lo_x) && std::isfinite(hi_x) && std::isfinite(lo_y) && std::isfinite(hi_y)
            && std::fabs(lo_x) < 1e9 && std::fabs(hi_x) < 1e9 && std::fabs(lo_y) < 1e9 && std::fabs(hi_y) < 1e9) {
            // margin: 2 px + a relative part for the rounding of the inverse on large coordinates
            const double mx = 2.0 + 1e-9 * (std::fabs(lo_x) + std::fabs(hi_x)), my = 2.0 + 1e-9 * (std::fabs(lo_y) + std::fabs(hi_y));
            p.bb_j0 = (int) std::floor(lo_x - mx), p.bb_j1 = (int) std::ceil(hi_x + mx);
            p.bb_i0 = (int) std::floor(lo_y - my), p.bb_i1 = (int) std::ceil(hi_y + my);
        }
    }
    int need = d->interp == MLTCUDA_INTERP_BL ? 2 : d->interp == MLTCUDA_INTERP_BC ? 4 : 1;
    if (!p.degenerate && (d->b_width < need || d->b_height < need)) {
        mltcuda_set_error("affine: B image %dx%d smaller than the %dx%d sampler footprint", d->b_width, d->b_height, need, need);
        return MLTCUDA_E_BADARG;
    }
    return 0;
}

extern "C" int mltcuda_affine(mltcuda_ctx *ctx, const mltcuda_affine_desc *d)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!d || !d->a_image || !d->b_image, "null image pointer");
    MC_BADARG(d->format != MLTCUDA_FMT_RGBA && d->format != MLTCUDA_FMT_RGBA64, "affine needs rgba or rgba64");
    MC_BADARG(d->interp < 0 || d->interp > 2, "bad interpolation");
    if (d->a_width <= 0 || d->a_height <= 0)
        return 0;
    AffK K;
    if (int e = mc_fill_affine_params(K.p, d))
        return e;
    K.a_image = (uint8_t *) d->a_image, K.b_image = (const uint8_t *) d->b_image;
    K.a_width = d->a_width, K.a_height = d->a_height;
    K.batch = d->batch > 0 ? d->batch : 1;
    K.a_bs = d->a_bstride, K.b_bs = d->b_bstride;
    if (K.batch > 1)
        MC_BADARG(!K.a_bs, "batch > 1 needs a_bstride");
    const int ppt = d->format == MLTCUDA_FMT_RGBA ? 4 : 2;
    long long total = (((long long) d->a_width * d->a_height + ppt - 1) / ppt) * K.batch;
    const bool rows_ok = !K.p.degenerate && d->a_width % ppt == 0 && ((uintptr_t) d->a_image & 15) == 0 && (K.a_bs & 15) == 0
                         && (d->a_height + 3) / 4 <= 65535 && K.batch <= 65535;
    if (rows_ok && d->format == MLTCUDA_FMT_RGBA)
        launch_affine_rows<uint8_t>(K, d->interp, ctx->stream);
    else if (rows_ok)
        launch_affine_rows<uint16_t>(K, d->interp, ctx->stream);
    else if (d->format == MLTCUDA_FMT_RGBA)
        k_affine<uint8_t><<<mc_blocks(total, 256), 256, 0, ctx->stream>>>(K);
    else
        k_affine<uint16_t><<<mc_blocks(total, 256), 256, 0, ctx->stream>>>(K);
    return mc_after_launch(ctx, "k_affine");
}
