// composite.cu -- K2: the composite_yuv row loop (reference
// src/modules/core/transition_composite.c:143-275,455-489 and
// src/modules/core/composite_line_yuv_sse2_simple.c:198-244) as one launch.
//
// Work item: one 16-byte-ALIGNED group of 8 destination pixels of one row (so A is read and written
// with 128-bit transactions whatever the rectangle's x offset); the B row, alpha_b and the luma map
// are fetched with aligned 128/64-bit loads + funnel shifts (fetch16/fetch8) because the rectangle
// shifts them by an arbitrary number of pixels relative to A.  Pixels of the group that fall
// outside [0,width) are masked.  The per-pixel arithmetic picks the SSE2 or the scalar formula
// exactly where the x86-64 reference does.  HBM-bound: 6 B/px (+1 alpha_b,
// +2 alpha_a r/w, +2 luma).
#include "common.cuh"
#include "pixel_ops.cuh"
#include <math.h>

namespace {

struct CompK
{
    CompositeParams p;
    uint8_t *dest;
    const uint8_t *src, *alpha_b;
    uint8_t *alpha_a;
    const uint16_t *luma;
    int rows;
    long stride_dest, stride_src, stride_aa, stride_ab;
    size_t dest_bs, src_bs, aa_bs, ab_bs, luma_bs;
    int batch;
    int cpr; // chunks per row upper bound
};

// One 16-byte-aligned group `c` of row `r` of frame `b`.  EDGE = false: whole groups only (all 8 pixels inside the
// row, one formula) -- the per-pixel code is compiled out and ragged groups are left to the EDGE = true pass.
// BLIT: P.blit (SSE2 cases 0/4), a separate instantiation so the blend cases keep their exact code.
template <bool BLIT, bool EDGE> __device__ __forceinline__ void composite_group(const CompK &K, int c, int r, int b)
{
    const CompositeParams &P = K.p;
    uintptr_t drow = (uintptr_t) (K.dest + b * K.dest_bs + r * K.stride_dest);
    int head = (int) ((drow & 15) >> 1);      // pixels between the aligned address and the row start
    int j0 = c * 8 - head;                    // row-relative index of the group's first pixel
    if (j0 >= P.width)
        return;
    int jlo = j0 < 0 ? 0 : j0, jhi = j0 + 8 > P.width ? P.width : j0 + 8; // valid [jlo, jhi)
    if (mc_composite_group_is_whole(P, j0, jlo, jhi) == EDGE)
        return; // the other pass owns this group
    uint4 *dptr = reinterpret_cast<uint4 *>(drow + 2 * (long) j0);        // 16-byte aligned
    // a whole group of a blit (SSE2 cases 0/4: memcpy + memset) overwrites A: do not read it
    const bool overwrite = BLIT && !EDGE;
    uint4 dv = make_uint4(0, 0, 0, 0);
    if (!overwrite)
        dv = ld16(dptr);
    uint32_t d[4] = {dv.x, dv.y, dv.z, dv.w};

    uintptr_t srow = (uintptr_t) (K.src + b * K.src_bs + r * K.stride_src);
    uint32_t s[4];
    fetch16(srow + 2 * (long) j0, srow + 2 * (long) jlo, srow + 2 * (long) jhi, s);

    uint32_t ab[2] = {0, 0}, aa[2] = {0, 0}, lum[4] = {0, 0, 0, 0};
    if (P.has_ab) {
        uintptr_t arow = (uintptr_t) (K.alpha_b + b * K.ab_bs + r * K.stride_ab);
        fetch8(arow + j0, arow + jlo, arow + jhi, ab);
    }
    uint8_t *aarow = nullptr;
    if (P.has_aa) {
        aarow = K.alpha_a + b * K.aa_bs + r * K.stride_aa;
        // read-modify-write: coherent byte/vector loads (no __ldg)
        uintptr_t a = (uintptr_t) aarow + j0;
        if (overwrite) {
        } else if (!EDGE && (a & 7) == 0) {
            uint2 v = ld8((const void *) a);
            aa[0] = v.x, aa[1] = v.y;
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (j0 + k >= jlo && j0 + k < jhi)
                    aa[k >> 2] |= (uint32_t) aarow[j0 + k] << ((k & 3) * 8);
        }
    }
    if (P.has_luma) {
        uintptr_t lrow = (uintptr_t) (K.luma + b * (K.luma_bs / 2) + r * K.stride_ab);
        fetch16(lrow + 2 * (long) j0, lrow + 2 * (long) jlo, lrow + 2 * (long) jhi, lum);
    }

    mc_composite_group8<true, EDGE>(P, j0, jlo, jhi, d, s, ab, aa, lum);
    const uint32_t *o = d, *oa = aa;
    if (!EDGE) {
        *dptr = make_uint4(o[0], o[1], o[2], o[3]);
    } else {
        uint16_t *d16 = reinterpret_cast<uint16_t *>(dptr);
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (j0 + k >= jlo && j0 + k < jhi)
                d16[k] = (uint16_t) (o[k >> 1] >> ((k & 1) * 16));
    }
    if (P.has_aa) {
        uintptr_t a = (uintptr_t) aarow + j0;
        if (!EDGE && (a & 7) == 0) {
            *reinterpret_cast<uint2 *>(a) = make_uint2(oa[0], oa[1]);
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (j0 + k >= jlo && j0 + k < jhi)
                    aarow[j0 + k] = (uint8_t) (oa[k >> 2] >> ((k & 3) * 8));
        }
    }
}

// interior pass: grid = (groups of a row, row, frame): no div/mod in the index math
template <bool BLIT> __global__ void __launch_bounds__(256, 8) k_composite_rows(CompK K)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= K.cpr)
        return;
    composite_group<BLIT, false>(K, c, blockIdx.y, blockIdx.z);
}

// Edge pass: the (at most three) groups of a row that are not whole -- the ragged first and last group of a
// rectangle that does not start / end on a 16-byte boundary of A, and the group that straddles the SSE2 / scalar
// boundary n8.  Each takes the per-pixel path (several hundred instructions); inside the interior
// kernel one such lane per row end kept its whole warp waiting (70 % of the instructions of a narrow rectangle at
// an odd x, profiles/r1l_weak_1).  Here every lane of a warp is a ragged group of a different row.
template <bool BLIT> __global__ void __launch_bounds__(128) k_composite_edges(CompK K)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= K.rows)
        return;
    const int b = blockIdx.z, e = blockIdx.y; // e: 0 first group, 1 the n8 straddler, 2 last group
    const uintptr_t drow = (uintptr_t) (K.dest + b * K.dest_bs + r * K.stride_dest);
    const int head = (int) ((drow & 15) >> 1);
    const int c_last = (head + K.p.width - 1) >> 3, c_mid = (head + K.p.n8) >> 3;
    int c;
    if (e == 0)
        c = 0;
    else if (e == 2)
        c = c_last;
    else
        c = c_mid;
    if ((e == 2 && c_last == 0) || (e == 1 && (c_mid == 0 || c_mid >= c_last)))
        return; // same group as an earlier e
    composite_group<BLIT, true>(K, c, r, b); // returns at once when the group is whole
}

} // namespace

void mc_fill_composite_params(CompositeParams &p, const mltcuda_composite_desc *d)
{
    p.width = d->width;
    p.weight = d->weight, p.softness = d->softness, p.step = d->step, p.op = d->op;
    p.has_ab = d->alpha_b != nullptr, p.has_aa = d->alpha_a != nullptr, p.has_luma = d->luma != nullptr;
    p.w8 = d->weight >> 8;
    p.soft_magic = mc_div_magic((uint32_t) d->softness);
    p.n8 = 0, p.blit = 0, p.premul = 0;
    // transition_composite.c:156-167: SSE2 routine only for composite_line_yuv itself (op over)
    if (d->op == MLTCUDA_OP_OVER && !d->scalar_only && !d->luma && d->width > 7) {
        p.n8 = d->width - d->width % 8;
        int cond = (p.has_aa ? 4 : 0) | (p.has_ab ? 2 : 0) | (p.w8 != 256 ? 1 : 0);
        p.blit = (cond == 0 || cond == 4);
        p.premul = cond & 1;
    }
}

extern "C" int mltcuda_composite(mltcuda_ctx *ctx, const mltcuda_composite_desc *d)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!d || !d->dest || !d->src, "null image pointer");
    MC_BADARG(d->op < 0 || d->op > 3, "bad operator");
    if (d->width <= 0 || d->rows <= 0)
        return 0;
    MC_BADARG(((uintptr_t) d->dest & 1) || ((uintptr_t) d->src & 1) || (d->stride_dest & 1) || (d->stride_src & 1),
              "yuv422 pointers/strides must be 2-byte aligned");
    CompK K;
    mc_fill_composite_params(K.p, d);
    K.dest = d->dest, K.src = d->src, K.alpha_b = d->alpha_b, K.alpha_a = d->alpha_a, K.luma = d->luma;
    K.rows = d->rows;
    K.stride_dest = d->stride_dest, K.stride_src = d->stride_src;
    K.stride_aa = d->stride_alpha_a, K.stride_ab = d->stride_alpha_b;
    K.batch = d->batch > 0 ? d->batch : 1;
    K.dest_bs = d->dest_bstride, K.src_bs = d->src_bstride, K.aa_bs = d->alpha_a_bstride;
    K.ab_bs = d->alpha_b_bstride, K.luma_bs = d->luma_bstride;
    if (K.batch > 1)
        MC_BADARG(!K.dest_bs, "batch > 1 needs dest_bstride");
    K.cpr = (d->width + 7 + 7) / 8; // up to 7 head pixels before the first composited pixel
    MC_BADARG(K.rows > 65535 || K.batch > 65535, "more than 65535 rows / frames in one launch");
    dim3 grd((K.cpr + 127) / 128, K.rows, K.batch);
    if (K.p.blit)
        k_composite_rows<true><<<grd, 128, 0, ctx->stream>>>(K);
    else
        k_composite_rows<false><<<grd, 128, 0, ctx->stream>>>(K);
    if (int e = mc_after_launch(ctx, "k_composite_rows"))
        return e;
    // Ragged groups exist unless every row starts on a 16-byte boundary of A and the width is a multiple of 8
    // (then n8 == width or 0 and every group is whole).
    const bool aligned = ((uintptr_t) d->dest & 15) == 0 && (d->stride_dest & 15) == 0 && (K.dest_bs & 15) == 0 && d->width % 8 == 0;
    if (aligned)
        return 0;
    dim3 egrd((K.rows + 127) / 128, 3, K.batch);
    if (K.p.blit)
        k_composite_edges<true><<<egrd, 128, 0, ctx->stream>>>(K);
    else
        k_composite_edges<false><<<egrd, 128, 0, ctx->stream>>>(K);
    return mc_after_launch(ctx, "k_composite_edges");
}

// Host restatement of the integer clipping in composite_yuv (transition_composite.c:345-452).
extern "C" int mltcuda_composite_setup(mltcuda_composite_desc *out, uint8_t *p_dest, int width_dest,
                                       int height_dest, const uint8_t *p_src, int width_src,
                                       int height_src, const uint8_t *alpha_b, uint8_t *alpha_a,
                                       const mltcuda_composite_geometry *g, int field,
                                       const uint16_t *p_luma, double softness)
{
    int x_src = -g->x_src, y_src = -g->y_src;
    int uneven_x_src = x_src % 2;
    int step = field > -1 ? 2 : 1;
    int stride_src = g->sw * 2, stride_dest = width_dest * 2;
    int x = (int) rint(g->x * width_dest / g->nw);
    int y = (int) rint(g->y * height_dest / g->nh);
    int uneven_x = x % 2;

    int op = out->op, scalar_only = out->scalar_only, batch = out->batch;
    size_t bs[5] = {out->dest_bstride, out->src_bstride, out->alpha_a_bstride, out->alpha_b_bstride, out->luma_bstride};
    *out = mltcuda_composite_desc();
    out->op = op, out->scalar_only = scalar_only, out->batch = batch;
    out->dest_bstride = bs[0], out->src_bstride = bs[1], out->alpha_a_bstride = bs[2];
    out->alpha_b_bstride = bs[3], out->luma_bstride = bs[4];
    out->softness = (int) ((1 << 16) * softness);
    out->weight = (int) (((1 << 16) * g->o + 50) / 100);
    out->step = (uint32_t) ((((1 << 16) - 1) * g->o + 50) / 100 * (1.0 + softness));

    if (width_src <= 0 || height_src <= 0 || y_src >= height_src || x_src >= width_src)
        return 1;
    if ((x < 0 && -x >= width_src) || (y < 0 && -y >= height_src))
        return 1;
    if (x_src > 0) {
        width_src -= x_src;
        if (width_src > g->w)
            width_src = (int) g->w;
    }
    if (y_src > 0) {
        height_src -= y_src;
        if (height_src > g->h)
            height_src = (int) g->h;
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
    out->dest = p_dest, out->src = p_src, out->alpha_b = alpha_b, out->alpha_a = alpha_a, out->luma = p_luma;
    out->width = width_src;
    out->rows = height_src > 0 ? (height_src + step - 1) / step : 0;
    out->stride_dest = stride_dest, out->stride_src = stride_src;
    out->stride_alpha_a = stride_dest / 2, out->stride_alpha_b = stride_src / 2;
    return (out->width <= 0 || out->rows <= 0) ? 1 : 0;
}
