// tractor.cu -- K5: fused multi-track pass.  Replaces the chain of full-frame read-modify-write
// passes (one per transition, plus a format conversion between neighbours that want different
// formats) that the reference's tractor / transition / convert plumbing produces
// (src/framework/mlt_tractor.c:348-424, mlt_transition.c:328-381, mlt_frame.c:501-541,839-998;
// SURVEY 3.4) by ONE kernel: each thread carries an 8-pixel group of the A frame in registers
// through the whole op list (composite rows, yuv422<->rgba conversions, affine sampling, opaque
// luma wipe) and writes the result once.  Every op is pointwise in A (B layers are gathered), and
// the group is pair-aligned, so all intermediate integer steps -- chroma pair averaging, clamps,
// alpha-plane creation -- are reproduced exactly: the output is byte-identical to running the ops
// as separate launches.  HBM traffic: A read once + written once + each layer's pixels once.
#include <stdlib.h>
#include "common.cuh"
#include "pixel_ops.cuh"

extern void mc_fill_composite_params(CompositeParams &p, const mltcuda_composite_desc *d);
extern int mc_fill_affine_params(AffineParams &p, const mltcuda_affine_desc *d);

namespace {

struct TCompOp
{
    CompositeParams cp;
    const uint8_t *src, *alpha_b;
    const uint16_t *luma;
    int x, y, rows, row_step;
    unsigned y_span, x_span; // rows / columns (+7) the rectangle spans in the frame
    int row_mask, row_shift; // field stepping: (row_step == 2) ? 1 : 0
    long stride_src, stride_ab;
    size_t src_bs, ab_bs, luma_bs;
};
struct TAffOp
{
    AffineParams ap;
    const uint8_t *b_image;
    size_t b_bs;
};
struct TLumaOp
{
    const uint8_t *src;
    const uint16_t *map;
    int luma_width, width_src, height_src, x_diff, y_diff;
    uint32_t i_softness, step;
    uint64_t soft_magic;
    size_t src_bs;
};
struct TOp
{
    int kind, to_format;
    union {
        TCompOp c;
        TAffOp a;
        TLumaOp l;
    };
};
struct TractorK
{
    int width, height, in_format, n_ops, batch;
    const uint8_t *in, *in_alpha;
    uint8_t *out, *out_alpha;
    size_t in_bs, ia_bs, out_bs, oa_bs;
    TOp ops[MLTCUDA_TRACTOR_MAX_OPS];
};

// MODE 1 (ONLY_BL): every affine op of the list samples bilinearly (the default "consumer.rescale"), so the
// nearest / bicubic samplers are compiled out -- they would otherwise set the register budget.
//
// SIG: 0 = interpret the op list at run time.  Otherwise the op list itself, 4 bits per slot
// (1 composite, 2 affine, 3 luma, 4 convert->rgba, 5 convert->yuv422): the loop is fully unrolled,
// every branch on the op kind folds away and every op parameter becomes an immediate constant-bank
// operand instead of an indexed LDC.  The host instantiates the common graph shapes (SIGS below)
// and falls back to SIG 0 for anything else; both run the same code, so results are identical.
__host__ __device__ constexpr int sig_len(uint32_t sig)
{
    int n = 0;
    while (sig) {
        n++;
        sig >>= 4;
    }
    return n;
}

// index of the first slot of SIG holding `code`, or -1
__host__ __device__ constexpr int sig_first(uint32_t sig, uint32_t code)
{
    for (int i = 0; sig; i++, sig >>= 4)
        if ((sig & 15) == code)
            return i;
    return -1;
}

#ifndef K5_MINB
#define K5_MINB 8
#endif
#ifndef K5_ROWS
#define K5_ROWS 4 // rows (= warps) per CTA
#endif
#ifndef K5_GEN_MINB
#define K5_GEN_MINB 8
#endif
// MODE: 1 = every affine op samples bilinearly, 0 = every affine op samples nearest, 2 = anything else (sampler chosen
// at run time).  Modes 0 and 1 compile exactly one sampler in and take the warp-transposed affine pass.
template <int MODE, uint32_t SIG>
__global__ void __launch_bounds__(32 * K5_ROWS, MODE != 2 ? (SIG ? K5_MINB : K5_GEN_MINB) * 4 / K5_ROWS : 2) k_tractor(const __grid_constant__ TractorK K)
{
    constexpr bool ONLY_BL = MODE != 2; // "one sampler" (name kept from when bilinear was the only such mode)
    // 2-D CTA: 32 groups (256 px) x 4 rows (128-thread CTAs balance the heavy affine tiles better than 256: +2.8%) -- no div/mod, and the B taps of neighbouring rows share L1
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int r = blockIdx.y * K5_ROWS + threadIdx.y;
    // blockDim.x == 32: a warp is 32 consecutive groups (256 px) of one row
    const bool valid = c < (K.width >> 3) && r < K.height;
    const bool fullwarp = __ballot_sync(0xffffffffu, valid) == 0xffffffffu;
    if (!valid)
        return;
    __shared__ uint32_t s_xpose[K5_ROWS][256]; // per-warp staging for the transposed affine pass
    const int b = blockIdx.z;
    const int col0 = c * 8;
    const size_t pix0 = (size_t) r * K.width + col0;

    uint32_t yuv[4] = {0, 0, 0, 0}, rgba[8], al[2] = {0xffffffffu, 0xffffffffu};
    int fmt = K.in_format;
    int has_alpha = 0;
    if (fmt == MLTCUDA_FMT_YUV422) {
        uint4 v = ldg16(K.in + b * K.in_bs + pix0 * 2);
        yuv[0] = v.x, yuv[1] = v.y, yuv[2] = v.z, yuv[3] = v.w;
        if (K.in_alpha) {
            uint2 a = ldg8(K.in_alpha + b * K.ia_bs + pix0);
            al[0] = a.x, al[1] = a.y, has_alpha = 1;
        }
    } else {
        const uint8_t *p = K.in + b * K.in_bs + pix0 * 4;
        uint4 v0 = ldg16(p), v1 = ldg16(p + 16);
        rgba[0] = v0.x, rgba[1] = v0.y, rgba[2] = v0.z, rgba[3] = v0.w;
        rgba[4] = v1.x, rgba[5] = v1.y, rgba[6] = v1.z, rgba[7] = v1.w;
    }

    // The first affine op's warp-level classification (row terms + the mapped ends of the warp's
    // 256-pixel span, all fp64 and independent of pixel data) is computed here, right behind the A
    // load, so it runs under that load's latency instead of after the first use of A.
    constexpr int AFF0 = sig_first(SIG, 2);
    AffineRow preR = {0.0, 0.0};
    double pre_xw0 = 0.0;
    bool pre_rej = false, pre_all = false;
    if (ONLY_BL && AFF0 >= 0 && fullwarp) {
        const AffineParams &P = K.ops[AFF0 < 0 ? 0 : AFF0].a.ap;
        pre_rej = mc_affine_bb_reject(P, r, blockIdx.x * 256, 256); // warp-uniform integer pre-test
        if (!pre_rej) {
            preR = mc_affine_row(P, r);
            pre_xw0 = P.lower_x + (double) (blockIdx.x * 256);
            double ax, ay, bx, by;
            mc_affine_map_x(P, preR, pre_xw0, ax, ay);
            mc_affine_map_x(P, preR, pre_xw0 + 255.0, bx, by);
            pre_rej = mc_affine_reject2(P, ax, ay, bx, by);
            pre_all = mc_affine_inside(P, ax, ay) && mc_affine_inside(P, bx, by);
        }
    }

    const int n_ops = SIG ? sig_len(SIG) : K.n_ops;
#pragma unroll
    for (int oi = 0; oi < n_ops; oi++) {
        const TOp &op = K.ops[oi];
        const int code = SIG ? (int) ((SIG >> (4 * oi)) & 15)
                             : (op.kind == MLTCUDA_TOP_CONVERT ? (op.to_format == MLTCUDA_FMT_RGBA ? 4 : 5) : op.kind);
        if (code == 1) {
            const TCompOp &C = op.c;
            // is this group inside the rectangle?  one unsigned compare per axis (y_span / x_span are set by the host)
            const int rrel = r - C.y;
            const int j0 = col0 - C.x;
            if ((unsigned) rrel >= C.y_span || (unsigned) (j0 + 7) >= C.x_span || (rrel & C.row_mask))
                continue;
            const int rr = rrel >> C.row_shift;
            int jlo = j0 < 0 ? 0 : j0, jhi = j0 + 8 > C.cp.width ? C.cp.width : j0 + 8;
            uint32_t s[4], ab[2] = {0, 0}, lum[4] = {0, 0, 0, 0};
            uintptr_t srow = (uintptr_t) (C.src + b * C.src_bs + rr * C.stride_src);
            fetch16(srow + 2 * (long) j0, srow + 2 * (long) jlo, srow + 2 * (long) jhi, s);
            if (C.cp.has_ab) {
                uintptr_t arow = (uintptr_t) (C.alpha_b + b * C.ab_bs + rr * C.stride_ab);
                fetch8(arow + j0, arow + jlo, arow + jhi, ab);
            }
            if (C.cp.has_luma) {
                uintptr_t lrow = (uintptr_t) (C.luma + b * (C.luma_bs / 2) + rr * C.stride_ab);
                fetch16(lrow + 2 * (long) j0, lrow + 2 * (long) jlo, lrow + 2 * (long) jhi, lum);
            }
            uint32_t na[2] = {al[0], al[1]};
            mc_composite_group8<false>(C.cp, j0, jlo, jhi, yuv, s, ab, na, lum);
            if (has_alpha)
                al[0] = na[0], al[1] = na[1];
        } else if (code == 4 || code == 5) {
            if (code == 4) { // filter_imageconvert.c:58-111
#pragma unroll
                for (int k = 0; k < 4; k++) // al[] is all 0xff while the frame has no alpha plane
                    mc_yuyv_to_rgba2w(yuv[k], al[k >> 1], (k & 1) * 2, (k & 1) * 2 + 1, rgba[2 * k], rgba[2 * k + 1]);
                fmt = MLTCUDA_FMT_RGBA, has_alpha = 0; // convert_image clears the alpha plane (:475-477)
            } else { // filter_imageconvert.c:144-184: creates the alpha plane
#pragma unroll
                for (int k = 0; k < 4; k++)
                    yuv[k] = mc_rgba2_to_yuyv(rgba[2 * k], rgba[2 * k + 1]);
                al[0] = __byte_perm(__byte_perm(rgba[0], rgba[1], 0x0073), __byte_perm(rgba[2], rgba[3], 0x0073), 0x5410);
                al[1] = __byte_perm(__byte_perm(rgba[4], rgba[5], 0x0073), __byte_perm(rgba[6], rgba[7], 0x0073), 0x5410);
                fmt = MLTCUDA_FMT_YUV422, has_alpha = 1;
            }
        } else if (code == 2) {
            const AffineParams &P = op.a.ap;
            const bool pre = ONLY_BL && AFF0 >= 0 && oi == AFF0;
            const AffineRow R = pre ? preR : mc_affine_row(P, r);
            if (ONLY_BL && fullwarp) {
                // Transposed pass.  With 8 consecutive pixels per thread the 32 lanes of a tap load
                // are 8 destination pixels apart -- up to 32 different L1 lines per LDG, which made the
                // taps the latency AND the L1-wavefront limit of the kernel (profiles/r1f).  Here the
                // warp's 256 pixels go through shared memory so that lane i samples pixels i, i+32, ...:
                // one tap load then touches 2-4 adjacent lines.  Same per-pixel arithmetic as below.
                __syncwarp();
                const int wcol0 = blockIdx.x * 256;
                double xw0;
                bool rej, all;
                if (pre) {
                    xw0 = pre_xw0, rej = pre_rej, all = pre_all;
                } else {
                    xw0 = P.lower_x + (double) wcol0;
                    double ax, ay, bx, by;
                    mc_affine_map_x(P, R, xw0, ax, ay);
                    mc_affine_map_x(P, R, xw0 + 255.0, bx, by);
                    rej = mc_affine_reject2(P, ax, ay, bx, by);
                    all = mc_affine_inside(P, ax, ay) && mc_affine_inside(P, bx, by);
                }
                if (rej)
                    continue; // warp-uniform
                const uint8_t *bimg = op.a.b_image + b * op.a.b_bs;
                asm volatile("" : "+l"(bimg));
                uint32_t *sm = s_xpose[threadIdx.y];
                reinterpret_cast<uint4 *>(sm)[2 * threadIdx.x] = make_uint4(rgba[0], rgba[1], rgba[2], rgba[3]);
                reinterpret_cast<uint4 *>(sm)[2 * threadIdx.x + 1] = make_uint4(rgba[4], rgba[5], rgba[6], rgba[7]);
                __syncwarp();
                const double xl = xw0 + (double) threadIdx.x;
#pragma unroll
                for (int h4 = 0; h4 < 8; h4 += 4) {
                    BlTaps<uint8_t> taps[4];
                    uint32_t ntap[4];
                    bool in[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        double dx, dy;
                        mc_affine_map_x(P, R, xl + (double) (32 * (h4 + k)), dx, dy);
                        in[k] = all || mc_affine_inside(P, dx, dy);
                        if (in[k]) {
                            if (MODE == 1)
                                taps[k] = mc_bl_fetch<uint8_t>(bimg, P.b_width, P.b_height, (float) dx, (float) dy);
                            else
                                ntap[k] = mc_nn_fetch<uint8_t>(bimg, P.b_width, (float) dx, (float) dy);
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        if (in[k]) {
                            const int p = 32 * (h4 + k) + threadIdx.x;
                            if (MODE == 1)
                                sm[p] = mc_bl_blend<uint8_t>(taps[k], P.mix, sm[p], P.b_alpha);
                            else
                                sm[p] = mc_nn_blend<uint8_t>(ntap[k], P.mix, sm[p], P.b_alpha);
                        }
                }
                __syncwarp();
                const uint4 q0 = reinterpret_cast<const uint4 *>(sm)[2 * threadIdx.x];
                const uint4 q1 = reinterpret_cast<const uint4 *>(sm)[2 * threadIdx.x + 1];
                rgba[0] = q0.x, rgba[1] = q0.y, rgba[2] = q0.z, rgba[3] = q0.w;
                rgba[4] = q1.x, rgba[5] = q1.y, rgba[6] = q1.z, rgba[7] = q1.w;
                continue;
            }
            const double x0 = P.lower_x + (double) col0; // + k below: see k_affine_rows
            double ax, ay, bx, by;
            mc_affine_map_x(P, R, x0, ax, ay);
            mc_affine_map_x(P, R, x0 + 7.0, bx, by);
            if (mc_affine_reject2(P, ax, ay, bx, by))
                continue;
            const uint8_t *bimg = op.a.b_image + b * op.a.b_bs;
            asm volatile("" : "+l"(bimg)); // keep the batch base in registers
            // both end pixels inside B => the whole group is (monotone mapping): no per-pixel tests
            if (MODE == 1 && mc_affine_inside(P, ax, ay) && mc_affine_inside(P, bx, by)) {
                // bilinear, whole group inside: taps of four pixels in flight before any blend
#pragma unroll
                for (int h4 = 0; h4 < 8; h4 += 4) {
                    BlTaps<uint8_t> taps[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        double dx = ax, dy = ay;
                        if (h4 + k == 7)
                            dx = bx, dy = by;
                        else if (h4 + k > 0)
                            mc_affine_map_x(P, R, x0 + (double) (h4 + k), dx, dy);
                        taps[k] = mc_bl_fetch<uint8_t>(bimg, P.b_width, P.b_height, (float) dx, (float) dy);
                    }
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        rgba[h4 + k] = mc_bl_blend<uint8_t>(taps[k], P.mix, rgba[h4 + k], P.b_alpha);
                }
            } else if (MODE != 1 && mc_affine_inside(P, ax, ay) && mc_affine_inside(P, bx, by)) {
#pragma unroll 1
                for (int k = 0; k < 8; k++) {
                    double dx = ax, dy = ay;
                    if (k == 7)
                        dx = bx, dy = by;
                    else if (k > 0)
                        mc_affine_map_x(P, R, x0 + (double) k, dx, dy);
                    if (MODE != 2)
                        rgba[k] = mc_affine_sample_packed<uint8_t, MODE == 1 ? 1 : 0>(P, bimg, dx, dy, rgba[k]);
                    else
                        rgba[k] = mc_affine_sample_packed_rt<uint8_t>(P, bimg, dx, dy, rgba[k]);
                }
            } else {
#pragma unroll 1
                for (int k = 0; k < 8; k++) {
                    double dx, dy;
                    mc_affine_map_x(P, R, x0 + (double) k, dx, dy);
                    if (mc_affine_inside(P, dx, dy)) {
                        if (MODE != 2)
                            rgba[k] = mc_affine_sample_packed<uint8_t, MODE == 1 ? 1 : 0>(P, bimg, dx, dy, rgba[k]);
                        else
                            rgba[k] = mc_affine_sample_packed_rt<uint8_t>(P, bimg, dx, dy, rgba[k]);
                    }
                }
            }
        } else if (code == 3) { // transition_luma.c:452-463, progressive
            const TLumaOp &L = op.l;
            if (r >= L.height_src || col0 >= L.width_src)
                continue;
            const uint8_t *srow = L.src + b * L.src_bs + (size_t) r * L.width_src * 2;
            uint32_t s[4];
            fetch16((uintptr_t) srow + 2 * col0, (uintptr_t) srow + 2 * col0,
                    (uintptr_t) srow + 2 * (col0 + 8 > L.width_src ? L.width_src : col0 + 8), s);
            const uint16_t *l = L.map + (long) ((r * L.y_diff) >> 16) * L.luma_width;
            if (col0 + 8 <= L.width_src) {
                uint32_t lum[8];
#pragma unroll
                for (int k = 0; k < 8; k++)
                    lum[k] = __ldg(l + (((uint32_t) (col0 + k) * (uint32_t) L.x_diff) >> 16));
                mc_luma_wipe_group8(s, yuv, lum, L.step, L.i_softness, L.soft_magic);
                continue;
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                int j = col0 + k;
                if (j < L.width_src) {
                    uint32_t dp = (yuv[k >> 1] >> ((k & 1) * 16)) & 0xffff;
                    uint32_t sp = (s[k >> 1] >> ((k & 1) * 16)) & 0xffff;
                    uint32_t w = __ldg(l + ((j * L.x_diff) >> 16));
                    uint32_t value = (uint32_t) mc_smoothstep((int32_t) w, (int32_t) (L.i_softness + w), L.step, L.soft_magic);
                    uint32_t qy = ((sp & 0xff) * value + (dp & 0xff) * ((1u << 16) - value)) >> 16;
                    uint32_t qc = ((sp >> 8) * value + (dp >> 8) * ((1u << 16) - value)) >> 16;
                    uint32_t res = (qy & 0xff) | ((qc & 0xff) << 8);
                    yuv[k >> 1] = (yuv[k >> 1] & ~(0xffffu << ((k & 1) * 16))) | (res << ((k & 1) * 16));
                }
            }
        }
    }

    if (fmt == MLTCUDA_FMT_YUV422) {
        *reinterpret_cast<uint4 *>(K.out + b * K.out_bs + pix0 * 2) = make_uint4(yuv[0], yuv[1], yuv[2], yuv[3]);
        if (K.out_alpha)
            *reinterpret_cast<uint2 *>(K.out_alpha + b * K.oa_bs + pix0) = make_uint2(al[0], al[1]);
    } else {
        uint8_t *p = K.out + b * K.out_bs + pix0 * 4;
        reinterpret_cast<uint4 *>(p)[0] = make_uint4(rgba[0], rgba[1], rgba[2], rgba[3]);
        reinterpret_cast<uint4 *>(p)[1] = make_uint4(rgba[4], rgba[5], rgba[6], rgba[7]);
    }
}

} // namespace

extern "C" int mltcuda_tractor(mltcuda_ctx *ctx, const mltcuda_tractor_desc *d)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!d || !d->in || !d->out, "null image pointer");
    MC_BADARG(d->n_ops < 0 || d->n_ops > MLTCUDA_TRACTOR_MAX_OPS, "too many ops");
    MC_BADARG(d->n_ops > 0 && !d->ops, "null op list");
    if (d->width <= 0 || d->height <= 0)
        return 0;
    if (d->width % 8) {
        mltcuda_set_error("fused tractor pass needs width %% 8 == 0 (got %d); run the ops unfused", d->width);
        return MLTCUDA_E_UNSUPPORTED;
    }
    MC_BADARG(d->in_format != MLTCUDA_FMT_YUV422 && d->in_format != MLTCUDA_FMT_RGBA, "in_format must be yuv422 or rgba");
    MC_BADARG(((uintptr_t) d->in | (uintptr_t) d->out | d->in_bstride | d->out_bstride) & 15, "frames must be 16-byte aligned");
    MC_BADARG(((uintptr_t) d->in_alpha | (uintptr_t) d->out_alpha | d->in_alpha_bstride | d->out_alpha_bstride) & 7,
              "alpha planes must be 8-byte aligned");
    static thread_local TractorK K; // ~3 KB, passed by value as a __grid_constant__
    K.width = d->width, K.height = d->height, K.in_format = d->in_format, K.n_ops = d->n_ops;
    K.batch = d->batch > 0 ? d->batch : 1;
    K.in = (const uint8_t *) d->in, K.in_alpha = d->in_alpha, K.out = (uint8_t *) d->out, K.out_alpha = d->out_alpha;
    K.in_bs = d->in_bstride, K.ia_bs = d->in_alpha_bstride, K.out_bs = d->out_bstride, K.oa_bs = d->out_alpha_bstride;
    if (K.batch > 1)
        MC_BADARG(!K.in_bs || !K.out_bs, "batch > 1 needs in/out bstride");

    int fmt = d->in_format, has_alpha = d->in_format == MLTCUDA_FMT_YUV422 && d->in_alpha != nullptr;
    bool only_bl = true, only_nn = true;
    for (int i = 0; i < d->n_ops; i++) {
        const mltcuda_tractor_op &o = d->ops[i];
        TOp &t = K.ops[i];
        t.kind = o.kind, t.to_format = o.to_format;
        if (o.kind == MLTCUDA_TOP_COMPOSITE) {
            MC_BADARG(fmt != MLTCUDA_FMT_YUV422, "composite op needs the running frame in yuv422");
            mltcuda_composite_desc cd = o.u.composite;
            MC_BADARG(!cd.src || cd.op < 0 || cd.op > 3, "bad composite op");
            cd.alpha_a = has_alpha ? (uint8_t *) 1 : nullptr; // only its presence matters
            mc_fill_composite_params(t.c.cp, &cd);
            long off = (long) (cd.dest - (uint8_t *) d->out);
            long rowb = (long) d->width * 2;
            MC_BADARG(off < 0 || off >= rowb * d->height, "composite dest must point into `out`");
            t.c.y = (int) (off / rowb), t.c.x = (int) ((off % rowb) / 2);
            MC_BADARG(cd.stride_dest != rowb && cd.stride_dest != 2 * rowb, "composite stride_dest must be 1 or 2 rows");
            t.c.row_step = (int) (cd.stride_dest / rowb);
            t.c.rows = cd.rows;
            t.c.row_shift = t.c.row_mask = t.c.row_step == 2 ? 1 : 0;
            t.c.y_span = (unsigned) ((cd.rows - 1) * t.c.row_step + 1);
            t.c.x_span = (unsigned) (cd.width + 7);
            t.c.src = cd.src, t.c.alpha_b = cd.alpha_b, t.c.luma = cd.luma;
            t.c.stride_src = cd.stride_src, t.c.stride_ab = cd.stride_alpha_b;
            t.c.src_bs = cd.src_bstride, t.c.ab_bs = cd.alpha_b_bstride, t.c.luma_bs = cd.luma_bstride;
        } else if (o.kind == MLTCUDA_TOP_CONVERT) {
            MC_BADARG(o.to_format != MLTCUDA_FMT_YUV422 && o.to_format != MLTCUDA_FMT_RGBA, "convert op: yuv422 or rgba");
            MC_BADARG(o.to_format == fmt, "convert op to the current format");
            fmt = o.to_format;
            has_alpha = fmt == MLTCUDA_FMT_YUV422;
        } else if (o.kind == MLTCUDA_TOP_AFFINE) {
            MC_BADARG(fmt != MLTCUDA_FMT_RGBA || o.u.affine.format != MLTCUDA_FMT_RGBA, "affine op needs the running frame in rgba");
            MC_BADARG(!o.u.affine.b_image, "null B image");
            if (int e = mc_fill_affine_params(t.a.ap, &o.u.affine))
                return e;
            if (t.a.ap.degenerate) {
                mltcuda_set_error("fused tractor pass: affine B image %dx%d below the 2x2 footprint; run the ops unfused",
                                  o.u.affine.b_width, o.u.affine.b_height);
                return MLTCUDA_E_UNSUPPORTED;
            }
            t.a.b_image = (const uint8_t *) o.u.affine.b_image, t.a.b_bs = o.u.affine.b_bstride;
            only_bl = only_bl && t.a.ap.interp == MLTCUDA_INTERP_BL;
            only_nn = only_nn && t.a.ap.interp == MLTCUDA_INTERP_NN;
        } else if (o.kind == MLTCUDA_TOP_LUMA) {
            const mltcuda_luma_desc &l = o.u.luma;
            MC_BADARG(fmt != MLTCUDA_FMT_YUV422 || l.format != MLTCUDA_FMT_YUV422 || !l.map || l.field_order >= 0,
                      "luma op: opaque progressive yuv422 wipe only");
            MC_BADARG(l.width <= 0 || l.height <= 0 || !l.src, "bad luma op");
            t.l.src = (const uint8_t *) l.src, t.l.map = l.map, t.l.luma_width = l.luma_width;
            t.l.width_src = l.width_src > d->width ? d->width : l.width_src;
            t.l.height_src = l.height_src > d->height ? d->height : l.height_src;
            t.l.x_diff = (l.luma_width << 16) / l.width, t.l.y_diff = (l.luma_height << 16) / l.height;
            t.l.i_softness = (uint32_t) (l.softness * (1 << 16));
            t.l.soft_magic = mc_div_magic(t.l.i_softness);
            float fpos = (l.pos + 0 * l.frame_delta * 0.5f) * (1.f + l.softness);
            t.l.step = (uint32_t) ((1 << 16) * fpos);
            t.l.src_bs = l.src_bstride;
        } else {
            mltcuda_set_error("unknown tractor op kind %d", o.kind);
            return MLTCUDA_E_BADARG;
        }
    }
    MC_BADARG(fmt != d->out_format, "out_format does not match the format the op list ends in");
    MC_BADARG((d->height + K5_ROWS - 1) / K5_ROWS > 65535 || K.batch > 65535, "frame too tall / batch too large for one launch");
    dim3 blk(32, K5_ROWS, 1), grd((d->width / 8 + 31) / 32, (d->height + K5_ROWS - 1) / K5_ROWS, K.batch);
    // op-list signature (see k_tractor): specialised instantiations for the common graph shapes
    uint32_t sig = 0;
    bool has_affine = false;
    for (int i = d->n_ops - 1; i >= 0; i--) {
        const mltcuda_tractor_op &o = d->ops[i];
        uint32_t code = o.kind == MLTCUDA_TOP_CONVERT ? (o.to_format == MLTCUDA_FMT_RGBA ? 4u : 5u) : (uint32_t) o.kind;
        sig = (sig << 4) | code;
        has_affine = has_affine || o.kind == MLTCUDA_TOP_AFFINE;
    }
    // MLTCUDA_TRACTOR_GENERIC=1: always the run-time-interpreted instantiation (what an unlisted graph shape gets);
    // used by tools/ktractor.py to time it against the specialised ones
    static const bool force_generic = [] {
        const char *e = getenv("MLTCUDA_TRACTOR_GENERIC");
        return e && *e && *e != '0';
    }();
    const int mode = only_bl ? 1 : (only_nn && has_affine) ? 0 : 2;
    if (force_generic || d->n_ops == 0)
        sig = 0;
#define MC_TRACTOR_SIG(M, S) \
    case S: \
        k_tractor<M, S><<<grd, blk, 0, ctx->stream>>>(K); \
        break;
    if (mode == 1) {
        switch (sig) {
            MC_TRACTOR_SIG(1, 0x1)     // composite
            MC_TRACTOR_SIG(1, 0x11)    // composite x2
            MC_TRACTOR_SIG(1, 0x111)   // composite x3
            MC_TRACTOR_SIG(1, 0x1111)  // composite x4
            MC_TRACTOR_SIG(1, 0x3)     // luma wipe
            MC_TRACTOR_SIG(1, 0x13)    // luma wipe, composite (a wipe between two clips under an overlay)
            MC_TRACTOR_SIG(1, 0x31)    // composite, luma wipe
            MC_TRACTOR_SIG(1, 0x24)    // ->rgba, affine (rgba consumer: cfg3)
            MC_TRACTOR_SIG(1, 0x524)   // ->rgba, affine, ->yuv422
            MC_TRACTOR_SIG(1, 0x5241)  // composite, ->rgba, affine, ->yuv422
            MC_TRACTOR_SIG(1, 0x1524)  // ->rgba, affine, ->yuv422, composite
            MC_TRACTOR_SIG(1, 0x15241) // cfg4: composite, ->rgba, affine, ->yuv422, composite
        default:
            k_tractor<1, 0><<<grd, blk, 0, ctx->stream>>>(K);
        }
    } else if (mode == 0) { // nearest-neighbour sampling (consumer.rescale=nearest: previews)
        switch (sig) {
            MC_TRACTOR_SIG(0, 0x24)
            MC_TRACTOR_SIG(0, 0x524)
            MC_TRACTOR_SIG(0, 0x15241)
        default:
            k_tractor<0, 0><<<grd, blk, 0, ctx->stream>>>(K);
        }
    } else {
        k_tractor<2, 0><<<grd, blk, 0, ctx->stream>>>(K);
    }
#undef MC_TRACTOR_SIG
    return mc_after_launch(ctx, "k_tractor");
}
