// luma.cu -- K3: luma wipe and dissolve (reference src/modules/core/transition_luma.c:63-292,
// 327-723) plus scale_luma (src/modules/core/transition_composite.c:497-520).
//
// yuv422: one thread = one 16-byte-aligned group of 8 destination pixels (A read-modify-write with
// 128-bit transactions, B through fetch16, the 16-bit map gathered through the read-only path with
// the reference's 16.16 fixed-point index).  rgba / rgba64: one thread = 16 bytes of A.
// The opaque (integer) formula is bit-exact; the translucent (float) formula follows the reference's
// operation order (-fmad=false), +-1 LSB against its -ffast-math build.  Whether the translucent
// path is taken is decided ON THE DEVICE by a flag kernel (is_opaque, :34-41) unless the caller
// forces it, so no host round trip sits between the scan and the blend.
// HBM-bound: 6 B/px yuv422 (+ map once), 12 B/px rgba, 24 B/px rgba64.
#include "common.cuh"
#include "pixel_ops.cuh"

extern void mc_fill_composite_params(CompositeParams &p, const mltcuda_composite_desc *d);

namespace {

// flag |= any(byte != 0xff) over n bytes with the given element stride/offset
__global__ void __launch_bounds__(256) k_flag_not_ff(const uint8_t *p, size_t n, int stride, int *flag)
{
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    size_t step = (size_t) gridDim.x * blockDim.x;
    int bad = 0;
    for (; i < n; i += step)
        bad |= p[i * stride] != 0xff;
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0)
        *flag = 1;
}
__global__ void __launch_bounds__(256) k_flag_not_ffff(const uint16_t *p, size_t n, int stride, int *flag)
{
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    size_t step = (size_t) gridDim.x * blockDim.x;
    int bad = 0;
    for (; i < n; i += step)
        bad |= p[i * stride] != 0xffff;
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0)
        *flag = 1;
}

struct LumaK
{
    uint8_t *dest;
    const uint8_t *src;
    uint8_t *alpha_dest, *alpha_src;
    const uint16_t *map;
    int luma_width;
    int width_src, height_src, width_dest; // after the min() of :376-377
    int x_diff, y_diff;
    uint32_t i_softness;
    uint64_t soft_magic; // mc_div_magic(i_softness)
    float softness, field_pos0, field_pos1;
    int field_count, invert, alpha_over;
    const int *flag; // translucent flag on device (may be NULL => forced value)
    int forced;
    int rows0;       // rows processed in field 0 (for the running alpha index)
    int batch;
    size_t dest_bs, src_bs, ad_bs, as_bs;
    int cpr;
    // dissolve
    float weight;
    CompositeParams cp;
};

__device__ __forceinline__ void luma_row_info(const LumaK &K, int i, int &field, int &y_off, int &ordinal)
{
    if (K.field_count == 1) {
        field = 0, y_off = i * K.y_diff, ordinal = i;
    } else {
        field = i & 1;
        int n = i >> 1; // n-th row of its field
        y_off = (field << 16) + n * K.y_diff;
        ordinal = field ? K.rows0 + n : n;
    }
}

// luma_composite_yuv422 (:327-473) and, with map == NULL, dissolve_yuv422 (:131-190)
template <bool DISSOLVE> __global__ void __launch_bounds__(256, 8) k_luma_yuv422(LumaK K)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x; // grid = (chunks of a row, row, frame)
    if (c >= K.cpr)
        return;
    const int i = blockIdx.y, b = blockIdx.z;
    const int translucent = K.flag ? *K.flag : K.forced;

    uintptr_t drow = (uintptr_t) (K.dest + b * K.dest_bs + (size_t) i * K.width_dest * 2);
    int head = (int) ((drow & 15) >> 1);
    int j0 = c * 8 - head;
    if (j0 >= K.width_src)
        return;
    int jlo = j0 < 0 ? 0 : j0, jhi = j0 + 8 > K.width_src ? K.width_src : j0 + 8;
    uint4 *dptr = reinterpret_cast<uint4 *>(drow + 2 * (long) j0);
    uint4 dv = ld16(dptr);
    uint32_t d[4] = {dv.x, dv.y, dv.z, dv.w}, s[4];
    uintptr_t srow = (uintptr_t) (K.src + b * K.src_bs + (size_t) i * K.width_src * 2);
    fetch16(srow + 2 * (long) j0, srow + 2 * (long) jlo, srow + 2 * (long) jhi, s);

    int field, y_off, ordinal;
    luma_row_info(K, i, field, y_off, ordinal);
    const uint16_t *l = DISSOLVE ? nullptr : K.map + (long) (y_off >> 16) * (K.luma_width * K.field_count);
    const float fpos = field ? K.field_pos1 : K.field_pos0;
    // running alpha pointers of the reference: one element per processed pixel (:446-449); for a
    // dissolve they advance by the plane width per row (:181-185, dissolve_slice :97-118)
    uint8_t *ad = K.alpha_dest ? K.alpha_dest + b * K.ad_bs
                                     + (DISSOLVE ? (size_t) i * (translucent && K.alpha_over ? K.width_src : K.width_dest)
                                                 : (size_t) ordinal * K.width_src)
                               : nullptr;
    uint8_t *as = K.alpha_src ? K.alpha_src + b * K.as_bs + (DISSOLVE ? (size_t) i * K.width_src : (size_t) ordinal * K.width_src)
                              : nullptr;
    uint32_t o[4];
    if (!DISSOLVE && !translucent && jlo == j0 && jhi == j0 + 8) {
        // opaque wipe on a whole group (:452-463): integer path without per-pixel predicates
        const uint32_t step = (uint32_t) ((1 << 16) * fpos);
        uint32_t lum[8];
        if (K.x_diff == 65536) { // map as wide as the frame: 8 consecutive entries = 16 bytes
            uint32_t lw[4];
            fetch16((uintptr_t) (l + j0), (uintptr_t) l, (uintptr_t) (l + K.luma_width), lw);
#pragma unroll
            for (int k = 0; k < 8; k++)
                lum[k] = (k & 1) ? lw[k >> 1] >> 16 : lw[k >> 1] & 0xffff;
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++)
                lum[k] = __ldg(l + (((uint32_t) (j0 + k) * (uint32_t) K.x_diff) >> 16));
        }
        mc_luma_wipe_group8(s, d, lum, step, K.i_softness, K.soft_magic);
        o[0] = d[0], o[1] = d[1], o[2] = d[2], o[3] = d[3];
        *dptr = make_uint4(o[0], o[1], o[2], o[3]);
        return;
    }
    if (DISSOLVE && !(translucent && K.alpha_over) && jlo == j0 && jhi == j0 + 8
        && ((((uintptr_t) (ad ? ad + j0 : nullptr)) | ((uintptr_t) (as ? as + j0 : nullptr))) & 7) == 0) {
        // cross-fade of a whole group = one composite_line_yuv group (:178-186): the composite
        // kernel's whole-group paths (SSE2 formula / scalar), alpha planes as 8-byte vectors
        uint32_t ab[2] = {0, 0}, aa[2] = {0, 0};
        const uint32_t lum0[4] = {0, 0, 0, 0};
        if (as) {
            const uint2 v = ld8(as + j0);
            ab[0] = v.x, ab[1] = v.y;
        }
        if (ad) {
            const uint2 v = ld8(ad + j0);
            aa[0] = v.x, aa[1] = v.y;
        }
        mc_composite_group8(K.cp, j0, jlo, jhi, d, s, ab, aa, lum0);
        *dptr = make_uint4(d[0], d[1], d[2], d[3]);
        if (ad)
            *reinterpret_cast<uint2 *>(ad + j0) = make_uint2(aa[0], aa[1]);
        return;
    }
    if (translucent && jlo == j0 && jhi == j0 + 8 && (!DISSOLVE || K.alpha_over)
        && ((((uintptr_t) (ad ? ad + j0 : nullptr)) | ((uintptr_t) (as ? as + j0 : nullptr))) & 7) == 0) {
        // translucent wipe (:417-451) / dissolve (composite_line_yuv_float :63-86) on a whole group
        // whose alpha bytes are 8-byte aligned: vector alpha loads / stores, the per-pixel float chain
        // on pixel pairs in packed fp32 -- the operations of the loop below, lane for lane
        const float2 unbias = make_float2(-8388608.0f, -8388608.0f);
        uint32_t adw[2] = {0xffffffffu, 0xffffffffu}, asw[2] = {0xffffffffu, 0xffffffffu}, aow[2];
        if (ad) {
            const uint2 v = ld8(ad + j0);
            adw[0] = v.x, adw[1] = v.y;
        }
        if (as) {
            const uint2 v = ld8(as + j0);
            asw[0] = v.x, asw[1] = v.y;
        }
        uint8_t *at = DISSOLVE ? ad : (K.invert && as) ? as : (!K.invert && ad) ? ad : nullptr;
        uint32_t o3[8];
#pragma unroll
        for (int kp = 0; kp < 4; kp++) { // pixels 2 kp, 2 kp + 1 = word kp
            const int aw_i = (2 * kp) >> 2, ab_i = (2 * kp) & 3;
            const float2 pa = f2_add(make_float2(__uint_as_float(__byte_perm(asw[aw_i], 0x4B000000u, 0x7540 | ab_i)),
                                                 __uint_as_float(__byte_perm(asw[aw_i], 0x4B000000u, 0x7540 | (ab_i + 1)))), unbias);
            const float2 qa = f2_add(make_float2(__uint_as_float(__byte_perm(adw[aw_i], 0x4B000000u, 0x7540 | ab_i)),
                                                 __uint_as_float(__byte_perm(adw[aw_i], 0x4B000000u, 0x7540 | (ab_i + 1)))), unbias);
            float2 value = f2_dup(K.weight);
            if (!DISSOLVE) {
                const uint32_t l0 = __ldg(l + (((j0 + 2 * kp) * K.x_diff) >> 16)), l1 = __ldg(l + (((j0 + 2 * kp + 1) * K.x_diff) >> 16));
                const float2 lf = f2_add(make_float2(__uint_as_float(0x4B000000u | l0), __uint_as_float(0x4B000000u | l1)), unbias);
                const float2 weight = mc_fdiv_const2<65535>(lf);
                value = mc_smoothstep_f2(weight, f2_add(f2_dup(K.softness), weight), fpos);
            }
            const float2 mix_a = mc_fdiv_const2<255>(f2_mul(f2_sub(f2_dup(1.0f), value), qa));
            float2 mix_b = mc_fdiv_const2<255>(f2_mul(value, pa));
            if (at) {
                const float2 mix2 = f2_add(mix_b, mix_a);
                if (DISSOLVE || K.alpha_over) {
                    const uint2 oa = mc_trunc_bits_pos2(f2_mul(f2_dup(255.0f), mix2));
                    o3[2 * kp] = oa.x, o3[2 * kp + 1] = oa.y;
                    const float2 q = mc_fdiv_fast_rn2(mix_b, mix2); // NaN where mix2 == 0: not selected
                    mix_b = make_float2(mix2.x != 0.f ? q.x : mix_b.x, mix2.y != 0.f ? q.y : mix_b.y);
                } else {
                    float2 tt = f2_lerp(at == as ? pa : qa, mix_b, 255.0f); // at[j] + 255 * mix_b, both rounded
                    tt.x = tt.x < 0.f ? 0.f : tt.x > 255.f ? 255.f : tt.x;
                    tt.y = tt.y < 0.f ? 0.f : tt.y > 255.f ? 255.f : tt.y;
                    o3[2 * kp] = mc_trunc_bits_pos(tt.x), o3[2 * kp + 1] = mc_trunc_bits_pos(tt.y);
                }
            }
            const float2 imix = f2_sub(f2_dup(1.0f), mix_b);
            const uint32_t sw = s[kp], dw = d[kp];
            const float2 py = f2_add(make_float2(__uint_as_float(__byte_perm(sw, 0x4B000000u, 0x7540)), __uint_as_float(__byte_perm(sw, 0x4B000000u, 0x7542))), unbias);
            const float2 pc = f2_add(make_float2(__uint_as_float(__byte_perm(sw, 0x4B000000u, 0x7541)), __uint_as_float(__byte_perm(sw, 0x4B000000u, 0x7543))), unbias);
            const float2 qy = f2_add(make_float2(__uint_as_float(__byte_perm(dw, 0x4B000000u, 0x7540)), __uint_as_float(__byte_perm(dw, 0x4B000000u, 0x7542))), unbias);
            const float2 qc = f2_add(make_float2(__uint_as_float(__byte_perm(dw, 0x4B000000u, 0x7541)), __uint_as_float(__byte_perm(dw, 0x4B000000u, 0x7543))), unbias);
            // p * mix_b + q * (1 - mix_b): both products rounded, then the sum
            const uint2 ty = mc_trunc_bits_pos2(f2_madd2r(f2_mul(py, mix_b), qy, imix));
            const uint2 tc = mc_trunc_bits_pos2(f2_madd2r(f2_mul(pc, mix_b), qc, imix));
            o[kp] = PxB<uint8_t>::pack(ty.x, tc.x, ty.y, tc.y);
        }
        *dptr = make_uint4(o[0], o[1], o[2], o[3]);
        if (at) {
            aow[0] = PxB<uint8_t>::pack(o3[0], o3[1], o3[2], o3[3]), aow[1] = PxB<uint8_t>::pack(o3[4], o3[5], o3[6], o3[7]);
            *reinterpret_cast<uint2 *>(at + j0) = make_uint2(aow[0], aow[1]);
        }
        return;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        int j = j0 + k;
        uint32_t dp = (d[k >> 1] >> ((k & 1) * 16)) & 0xffff;
        uint32_t sp = (s[k >> 1] >> ((k & 1) * 16)) & 0xffff;
        uint32_t res = dp;
        if (j >= jlo && j < jhi) {
            uint32_t qy = dp & 0xff, qc = dp >> 8, py = sp & 0xff, pc = sp >> 8;
            if (DISSOLVE) {
                if (translucent && K.alpha_over) { // composite_line_yuv_float :63-86
                    float mix_a = mc_fdiv_const<255>((1.0f - K.weight) * (float) (ad ? ad[j] : 255));
                    float mix_b = mc_fdiv_const<255>(K.weight * (float) (as ? as[j] : 255));
                    if (ad) {
                        float mix2 = mix_b + mix_a;
                        ad[j] = (uint8_t) mc_f2u_pos(255 * mix2, 0xff);
                        if (mix2 != 0.f)
                            mix_b = mc_fdiv_fast_rn(mix_b, mix2);
                    }
                    qy = mc_f2u_pos(py * mix_b + qy * (1.f - mix_b), 0xff);
                    qc = mc_f2u_pos(pc * mix_b + qc * (1.f - mix_b), 0xff);
                    res = qy | (qc << 8);
                } else { // composite_line_yuv(..., mix, NULL, 0, 0) :178-186
                    uint32_t aak = ad ? ad[j] : 0;
                    res = mc_composite_px(K.cp, j, dp, sp, as ? as[j] : 255, aak, 0);
                    if (ad)
                        ad[j] = (uint8_t) aak;
                }
            } else if (translucent) { // :417-451
                float weight = mc_fdiv_const<65535>((float) __ldg(l + ((j * K.x_diff) >> 16)));
                float value = mc_smoothstep_f(weight, K.softness + weight, fpos);
                float mix_a = mc_fdiv_const<255>((1.0f - value) * (float) (ad ? ad[j] : 255));
                float mix_b = mc_fdiv_const<255>(value * (float) (as ? as[j] : 255));
                uint8_t *at = (K.invert && as) ? as : (!K.invert && ad) ? ad : nullptr;
                if (at) {
                    float mix2 = mix_b + mix_a;
                    if (K.alpha_over) {
                        at[j] = (uint8_t) mc_f2u_pos(255 * mix2, 0xff);
                        if (mix2 != 0.f)
                            mix_b = mc_fdiv_fast_rn(mix_b, mix2);
                    } else {
                        float tt = (float) at[j] + 255 * mix_b;
                        tt = tt < 0.f ? 0.f : tt > 255.f ? 255.f : tt;
                        at[j] = (uint8_t) mc_f2u_pos(tt, 0xff);
                    }
                }
                qy = mc_f2u_pos(py * mix_b + qy * (1.f - mix_b), 0xff);
                qc = mc_f2u_pos(pc * mix_b + qc * (1.f - mix_b), 0xff);
                res = qy | (qc << 8);
            } else { // :452-463
                uint32_t weight = __ldg(l + ((j * K.x_diff) >> 16));
                uint32_t value = (uint32_t) mc_smoothstep((int32_t) weight, (int32_t) (K.i_softness + weight),
                                                          (uint32_t) ((1 << 16) * fpos), K.soft_magic);
                qy = (py * value + qy * ((1u << 16) - value)) >> 16;
                qc = (pc * value + qc * ((1u << 16) - value)) >> 16;
                res = (qy & 0xff) | ((qc & 0xff) << 8);
            }
        }
        if (k & 1)
            o[k >> 1] |= res << 16;
        else
            o[k >> 1] = res;
    }
    if (jlo == j0 && jhi == j0 + 8) {
        *dptr = make_uint4(o[0], o[1], o[2], o[3]);
    } else {
        uint16_t *d16 = reinterpret_cast<uint16_t *>(dptr);
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (j0 + k >= jlo && j0 + k < jhi)
                d16[k] = (uint16_t) (o[k >> 1] >> ((k & 1) * 16));
    }
}

// luma_composite_rgba32/_rgba64 (:475-723) and dissolve_slice_rgba32/_rgba64 (:192-265).
// T = uint8_t: 4 px per thread; uint16_t: 2 px per thread (16 bytes of A either way).
// WHOLE (row width a multiple of the group): every group is full and inside one row, so only the
// straight-line whole-group paths are instantiated (fewer registers: 8 CTAs/SM).
template <typename T, bool DISSOLVE, bool WHOLE>
__global__ void __launch_bounds__(256, (WHOLE || sizeof(T) == 2) ? 8 : 6) k_luma_rgba(LumaK K)
{
    constexpr int PPT = sizeof(T) == 1 ? 4 : 2;
    const float maxv = PxTraits<T>::maxv;
    constexpr int MAXI = sizeof(T) == 1 ? 255 : 65535;
    const uint32_t mask = PxTraits<T>::mask;
    const int w = K.width_src, h = K.height_src;
    const long long npx = (long long) w * h;
    const long long cpf = (npx + PPT - 1) / PPT;
    // K.cpr > 0 (row width a multiple of PPT): 2-D CTAs of 32 groups x 8 rows, no division.
    // Otherwise grid = (groups of a frame, 1, frame); a frame has < 2^31 pixels: 32-bit index math.
    unsigned gid;
    int i, j;
    if (WHOLE || K.cpr > 0) {
        const int g = blockIdx.x * 32 + threadIdx.x;
        i = blockIdx.y * 8 + threadIdx.y;
        if (g >= K.cpr || i >= h)
            return;
        j = g * PPT;
        gid = (unsigned) i * (unsigned) K.cpr + (unsigned) g;
    } else {
        gid = blockIdx.x * blockDim.x + threadIdx.x;
        if (gid >= (unsigned) cpf)
            return;
        i = (int) (gid * PPT / (unsigned) w), j = (int) (gid * PPT - (unsigned) i * (unsigned) w);
    }
    const int b = blockIdx.z;
    const unsigned p0 = gid * PPT;
    const int translucent = K.flag ? *K.flag : K.forced;
    T *q = reinterpret_cast<T *>(K.dest + b * K.dest_bs) + (size_t) p0 * 4;
    const T *p = reinterpret_cast<const T *>(K.src + b * K.src_bs) + (size_t) p0 * 4;
    uint32_t qv[PPT][4], pv[PPT][4];
    const bool full = WHOLE || p0 + PPT <= npx;
    if (WHOLE) {
    } else if (full) {
        uint4 a = ld16(q), s = ldg16(p);
        uint32_t aw[4] = {a.x, a.y, a.z, a.w}, sw[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
        for (int k = 0; k < PPT; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                if (sizeof(T) == 1)
                    qv[k][c] = (aw[k] >> (8 * c)) & 0xff, pv[k][c] = (sw[k] >> (8 * c)) & 0xff;
                else
                    qv[k][c] = (aw[2 * k + (c >> 1)] >> (16 * (c & 1))) & 0xffff,
                    pv[k][c] = (sw[2 * k + (c >> 1)] >> (16 * (c & 1))) & 0xffff;
            }
    } else {
#pragma unroll
        for (int k = 0; k < PPT; k++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                bool ok = p0 + k < npx;
                qv[k][c] = ok ? q[4 * k + c] : 0, pv[k][c] = ok ? p[4 * k + c] : 0;
            }
    }
    int field = 0, y_off = 0, ordinal = 0;
    const uint16_t *lrow = nullptr;
    float fpos = 0.f;
    if (!DISSOLVE && !translucent && full && (WHOLE || j + PPT <= w)) {
        // opaque wipe, whole group inside one row (:576-590, :693-707): RGB mixed in integers, alpha kept
        luma_row_info(K, i, field, y_off, ordinal);
        fpos = field ? K.field_pos1 : K.field_pos0;
        lrow = K.map + (long) (y_off >> 16) * (K.luma_width * K.field_count);
        const uint32_t step = (uint32_t) ((1 << 16) * fpos);
        uint32_t x[PPT], v[PPT];
        bool band = false;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            x[k] = step - __ldg(lrow + (((uint32_t) (j + k) * (uint32_t) K.x_diff) >> 16));
            band |= x[k] < K.i_softness;
        }
        uint4 a = ld16(q), s = ldg16(p);
        uint32_t aw[4] = {a.x, a.y, a.z, a.w}, sw[4] = {s.x, s.y, s.z, s.w};
        if (!band) { // every pixel fully A or fully B (alpha stays A's)
#pragma unroll
            for (int k = 0; k < PPT; k++) {
                if ((int32_t) x[k] >= 0) {
                    if (sizeof(T) == 1)
                        aw[k] = __byte_perm(sw[k], aw[k], 0x7210);
                    else
                        aw[2 * k] = sw[2 * k], aw[2 * k + 1] = __byte_perm(sw[2 * k + 1], aw[2 * k + 1], 0x7610);
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < PPT; k++) {
                const uint32_t t = mc_div_by(x[k] << 16, K.i_softness, K.soft_magic);
                const uint32_t poly = (((t * t) >> 16) * ((3u << 16) - (2 * t))) >> 16;
                v[k] = x[k] < K.i_softness ? poly : ((int32_t) x[k] < 0 ? 0u : 0x10000u);
                const uint32_t iv = 65536u - v[k];
                if (sizeof(T) == 1) { // byte 2 of each sum is the mixed channel
                    const uint32_t c0 = (sw[k] & 0xff) * v[k] + (aw[k] & 0xff) * iv;
                    const uint32_t c1 = __byte_perm(sw[k], 0, 0x4441) * v[k] + __byte_perm(aw[k], 0, 0x4441) * iv;
                    const uint32_t c2 = __byte_perm(sw[k], 0, 0x4442) * v[k] + __byte_perm(aw[k], 0, 0x4442) * iv;
                    aw[k] = __byte_perm(__byte_perm(c0, c1, 0x0062), __byte_perm(c2, aw[k], 0x0072), 0x5410);
                } else { // upper half of each sum
                    const uint32_t c0 = (sw[2 * k] & 0xffff) * v[k] + (aw[2 * k] & 0xffff) * iv;
                    const uint32_t c1 = (sw[2 * k] >> 16) * v[k] + (aw[2 * k] >> 16) * iv;
                    const uint32_t c2 = (sw[2 * k + 1] & 0xffff) * v[k] + (aw[2 * k + 1] & 0xffff) * iv;
                    aw[2 * k] = __byte_perm(c0, c1, 0x7632);
                    aw[2 * k + 1] = __byte_perm(c2, aw[2 * k + 1], 0x7632);
                }
            }
        }
        *reinterpret_cast<uint4 *>(q) = make_uint4(aw[0], aw[1], aw[2], aw[3]);
        return;
    }
    if (WHOLE || (full && (DISSOLVE || (translucent && j + PPT <= w)))) {
        // translucent wipe / dissolve, whole group (:192-265, :592-640, :709-757): the float formula of
        // the loop below, straight-line, with byte -> float by PRMT and truncation off the XU pipe
        const float2 unbias = make_float2(-8388608.0f, -8388608.0f);
        if (!DISSOLVE) {
            luma_row_info(K, i, field, y_off, ordinal);
            fpos = field ? K.field_pos1 : K.field_pos0;
            lrow = K.map + (long) (y_off >> 16) * (K.luma_width * K.field_count);
        }
        uint4 a = ld16(q), s = ldg16(p);
        uint32_t aw[4] = {a.x, a.y, a.z, a.w}, sw[4] = {s.x, s.y, s.z, s.w};
        // the per-pixel scalar chain (map weight, smoothstep, mix_a, mix_b, mix2, mix_b / mix2) runs on
        // two pixels at once in packed fp32 (FMUL2 / FFMA2): same operations per lane
#pragma unroll
        for (int k = 0; k < PPT; k += 2) {
            // alpha of pixels k, k + 1: the top byte / half of each pixel
            float2 pa, qa;
            if constexpr (sizeof(T) == 1) {
                pa = make_float2(__uint_as_float(__byte_perm(sw[k], 0x4B000000u, 0x7543)), __uint_as_float(__byte_perm(sw[k + 1], 0x4B000000u, 0x7543)));
                qa = make_float2(__uint_as_float(__byte_perm(aw[k], 0x4B000000u, 0x7543)), __uint_as_float(__byte_perm(aw[k + 1], 0x4B000000u, 0x7543)));
            } else {
                pa = make_float2(__uint_as_float(__byte_perm(sw[2 * k + 1], 0x4B000000u, 0x7532)), __uint_as_float(__byte_perm(sw[2 * k + 3], 0x4B000000u, 0x7532)));
                qa = make_float2(__uint_as_float(__byte_perm(aw[2 * k + 1], 0x4B000000u, 0x7532)), __uint_as_float(__byte_perm(aw[2 * k + 3], 0x4B000000u, 0x7532)));
            }
            pa = f2_add(pa, unbias), qa = f2_add(qa, unbias);
            float2 value = f2_dup(K.weight);
            if (!DISSOLVE) {
                const uint32_t l0 = __ldg(lrow + (((uint32_t) (j + k) * (uint32_t) K.x_diff) >> 16));
                const uint32_t l1 = __ldg(lrow + (((uint32_t) (j + k + 1) * (uint32_t) K.x_diff) >> 16));
                const float2 lf = f2_add(make_float2(__uint_as_float(0x4B000000u | l0), __uint_as_float(0x4B000000u | l1)), unbias);
                const float2 weight = mc_fdiv_const2<65535>(lf);
                value = mc_smoothstep_f2(weight, f2_add(f2_dup(K.softness), weight), fpos);
            }
            const float2 mix_a = mc_fdiv_const2<MAXI>(f2_mul(f2_sub(f2_dup(1.0f), value), qa));
            float2 mix_b = mc_fdiv_const2<MAXI>(f2_mul(value, pa));
            const float2 mix2 = f2_add(mix_b, mix_a);
            uint32_t o3[2];
            if (DISSOLVE || K.alpha_over) {
                const uint2 oa = mc_trunc_bits_pos2(f2_mul(f2_dup(maxv), mix2));
                o3[0] = oa.x, o3[1] = oa.y;
                const float2 d = mc_fdiv_fast_rn2(mix_b, mix2); // NaN where mix2 == 0: not selected
                mix_b = make_float2(mix2.x != 0.f ? d.x : mix_b.x, mix2.y != 0.f ? d.y : mix_b.y);
            } else {
                float2 tt = f2_lerp(qa, mix_b, maxv); // qa + maxv * mix_b, product and sum both rounded
                tt.x = tt.x < 0.f ? 0.f : tt.x > maxv ? maxv : tt.x;
                tt.y = tt.y < 0.f ? 0.f : tt.y > maxv ? maxv : tt.y;
                o3[0] = mc_trunc_bits_pos(tt.x), o3[1] = mc_trunc_bits_pos(tt.y);
            }
            const float2 imix = f2_sub(f2_dup(1.0f), mix_b);
#pragma unroll
            for (int h = 0; h < 2; h++) {
                PxB<T> P, Q;
                if constexpr (sizeof(T) == 1) {
                    P.from_word(sw[k + h]), Q.from_word(aw[k + h]);
                } else {
                    P.from_word(make_uint2(sw[2 * (k + h)], sw[2 * (k + h) + 1])), Q.from_word(make_uint2(aw[2 * (k + h)], aw[2 * (k + h) + 1]));
                }
                const float mb = h ? mix_b.y : mix_b.x, im = h ? imix.y : imix.x;
                const float2 p_lo = f2_add(P.lo, unbias), q_lo = f2_add(Q.lo, unbias);
                const float p_b = P.hi.x - 8388608.0f, q_b = Q.hi.x - 8388608.0f;
                // p * mix_b + q * (1 - mix_b): both products rounded, then the sum
                const float2 o_lo = f2_lerp(f2_mul(p_lo, f2_dup(mb)), q_lo, im);
                const float o_b = p_b * mb + q_b * im;
                // mixes of non-negative values with weights in [0, 1]: never negative, never NaN
                const uint2 t_lo = mc_trunc_bits_pos2(o_lo);
                if constexpr (sizeof(T) == 1) {
                    aw[k + h] = PxB<T>::pack(t_lo.x, t_lo.y, mc_trunc_bits_pos(o_b), o3[h]);
                } else {
                    const uint2 o = PxB<T>::pack(t_lo.x, t_lo.y, mc_trunc_bits_pos(o_b), o3[h]);
                    aw[2 * (k + h)] = o.x, aw[2 * (k + h) + 1] = o.y;
                }
            }
        }
        *reinterpret_cast<uint4 *>(q) = make_uint4(aw[0], aw[1], aw[2], aw[3]);
        return;
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        if (p0 + k < npx) {
            if (DISSOLVE) {
                float mix_a = mc_fdiv_const<MAXI>((1.0f - K.weight) * (float) qv[k][3]);
                float mix_b = mc_fdiv_const<MAXI>(K.weight * (float) pv[k][3]);
                float mix2 = mix_b + mix_a;
                if (mix2 != 0.f)
                    mix_b = mc_fdiv_fast_rn(mix_b, mix2);
#pragma unroll
                for (int c = 0; c < 3; c++)
                    qv[k][c] = mc_f2u_pos(pv[k][c] * mix_b + qv[k][c] * (1.f - mix_b), mask);
                qv[k][3] = mc_f2u_pos(maxv * mix2, mask);
            } else {
                if (k == 0 || j == 0) { // row-dependent terms: once per thread unless the group wraps
                    luma_row_info(K, i, field, y_off, ordinal);
                    fpos = field ? K.field_pos1 : K.field_pos0;
                    lrow = K.map + (long) (y_off >> 16) * (K.luma_width * K.field_count);
                }
                uint32_t lw = __ldg(lrow + ((j * K.x_diff) >> 16));
                if (translucent) {
                    float weight = mc_fdiv_const<65535>((float) lw);
                    float value = mc_smoothstep_f(weight, K.softness + weight, fpos);
                    float mix_a = mc_fdiv_const<MAXI>((1.0f - value) * (float) qv[k][3]);
                    float mix_b = mc_fdiv_const<MAXI>(value * (float) pv[k][3]);
                    float mix2 = mix_b + mix_a;
                    if (K.alpha_over) {
                        qv[k][3] = mc_f2u_pos(maxv * mix2, mask);
                        if (mix2 != 0.f)
                            mix_b = mc_fdiv_fast_rn(mix_b, mix2);
                    } else {
                        float tt = (float) qv[k][3] + maxv * mix_b;
                        tt = tt < 0.f ? 0.f : tt > maxv ? maxv : tt;
                        qv[k][3] = mc_f2u_pos(tt, mask);
                    }
#pragma unroll
                    for (int c = 0; c < 3; c++)
                        qv[k][c] = mc_f2u_pos(pv[k][c] * mix_b + qv[k][c] * (1.f - mix_b), mask);
                } else {
                    uint32_t value = (uint32_t) mc_smoothstep((int32_t) lw, (int32_t) (K.i_softness + lw),
                                                              (uint32_t) ((1 << 16) * fpos), K.soft_magic);
#pragma unroll
                    for (int c = 0; c < 3; c++)
                        qv[k][c] = ((pv[k][c] * value + qv[k][c] * ((1u << 16) - value)) >> 16) & mask;
                }
            }
        }
        if (++j == w)
            j = 0, i++;
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

__global__ void __launch_bounds__(256) k_scale_luma(uint16_t *dst, int dw, int dh, const uint16_t *src, int sw,
                                                    int x_step, int y_step, int invert_xor)
{
    long long gid = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (long long) dw * dh)
        return;
    int j = (int) (gid % dw), i = (int) (gid / dw);
    dst[gid] = __ldg(src + (long) ((i * y_step) >> 16) * sw + ((j * x_step) >> 16)) ^ (uint16_t) invert_xor;
}

} // namespace

extern "C" int mltcuda_scale_luma(mltcuda_ctx *ctx, uint16_t *dst, int dw, int dh, const uint16_t *src,
                                  int sw, int sh, int invert_xor)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || !src || dw <= 0 || dh <= 0 || sw <= 0 || sh <= 0, "bad argument");
    k_scale_luma<<<mc_blocks((long long) dw * dh, 256), 256, 0, ctx->stream>>>(dst, dw, dh, src, sw, (sw << 16) / dw,
                                                                            (sh << 16) / dh, invert_xor);
    return mc_after_launch(ctx, "k_scale_luma");
}

extern "C" int mltcuda_luma(mltcuda_ctx *ctx, const mltcuda_luma_desc *d)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!d || !d->dest || !d->src, "null image pointer");
    const int fmt = d->format;
    MC_BADARG(fmt != MLTCUDA_FMT_YUV422 && fmt != MLTCUDA_FMT_RGBA && fmt != MLTCUDA_FMT_RGBA64, "bad format");
    if (d->width == 0 || d->height == 0) // :369-370
        return 0;
    LumaK K = {};
    K.dest = (uint8_t *) d->dest, K.src = (const uint8_t *) d->src;
    K.alpha_dest = d->alpha_dest, K.alpha_src = d->alpha_src;
    K.map = d->map, K.luma_width = d->luma_width;
    K.batch = d->batch > 0 ? d->batch : 1;
    K.dest_bs = d->dest_bstride, K.src_bs = d->src_bstride, K.ad_bs = d->alpha_dest_bstride, K.as_bs = d->alpha_src_bstride;
    if (K.batch > 1)
        MC_BADARG(!K.dest_bs || !K.src_bs, "batch > 1 needs dest/src bstride");
    K.invert = d->invert, K.alpha_over = d->alpha_over;
    K.softness = d->softness;
    K.i_softness = (uint32_t) (d->softness * (1 << 16));
    K.soft_magic = mc_div_magic(K.i_softness);
    K.field_count = d->field_order < 0 ? 1 : 2;
    K.field_pos0 = (d->pos + ((d->field_order == 0 ? 1 : 0) * d->frame_delta * 0.5f)) * (1.f + d->softness);
    K.field_pos1 = (d->pos + ((d->field_order == 0 ? 0 : 1) * d->frame_delta * 0.5f)) * (1.f + d->softness);
    const bool dissolve = d->map == nullptr;
    K.weight = d->pos;

    // translucency flag
    int *flag = nullptr;
    K.forced = d->translucent > 0;
    if (d->translucent < 0) {
        MC_BADARG(K.batch != 1, "device-side opacity scan supports batch == 1");
        flag = (int *) mltcuda_malloc(ctx, sizeof(int));
        if (!flag)
            return (int) cudaErrorMemoryAllocation;
        MC_CHECK(cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream));
        const int grid = ctx->sm_count * 8;
        if (fmt == MLTCUDA_FMT_YUV422) {
            if (d->alpha_dest) {
                k_flag_not_ff<<<grid, 256, 0, ctx->stream>>>(d->alpha_dest, (size_t) d->width_dest * d->height_dest, 1, flag);
                mc_after_launch(ctx, "k_flag_not_ff");
            }
            if (d->alpha_src) {
                k_flag_not_ff<<<grid, 256, 0, ctx->stream>>>(d->alpha_src, (size_t) d->width_src * d->height_src, 1, flag);
                mc_after_launch(ctx, "k_flag_not_ff");
            }
        } else if (fmt == MLTCUDA_FMT_RGBA) {
            k_flag_not_ff<<<grid, 256, 0, ctx->stream>>>((const uint8_t *) d->dest + 3, (size_t) d->width_dest * d->height_dest, 4, flag);
            k_flag_not_ff<<<grid, 256, 0, ctx->stream>>>((const uint8_t *) d->src + 3, (size_t) d->width_src * d->height_src, 4, flag);
            mc_after_launch(ctx, "k_flag_not_ff");
            mc_after_launch(ctx, "k_flag_not_ff");
        } else {
            k_flag_not_ffff<<<grid, 256, 0, ctx->stream>>>((const uint16_t *) d->dest + 3, (size_t) d->width_dest * d->height_dest, 4, flag);
            k_flag_not_ffff<<<grid, 256, 0, ctx->stream>>>((const uint16_t *) d->src + 3, (size_t) d->width_src * d->height_src, 4, flag);
            mc_after_launch(ctx, "k_flag_not_ffff");
            mc_after_launch(ctx, "k_flag_not_ffff");
        }
        K.flag = flag;
    }

    int rc = 0;
    if (fmt == MLTCUDA_FMT_YUV422) {
        K.width_dest = d->width_dest;
        K.width_src = d->width_src > d->width_dest ? d->width_dest : d->width_src;
        K.height_src = d->height_src > d->height_dest ? d->height_dest : d->height_src;
        if (dissolve) {
            // dissolve_yuv422 :131-190: mix = weight * (1<<16), composite_line_yuv per row.
            mltcuda_composite_desc cd = {};
            cd.width = K.width_src, cd.weight = (int) (d->pos * (1 << 16));
            cd.alpha_b = d->alpha_src, cd.alpha_a = d->alpha_dest;
            mc_fill_composite_params(K.cp, &cd);
        } else {
            MC_BADARG(d->luma_width <= 0 || d->luma_height <= 0, "bad luma map size");
            K.x_diff = (d->luma_width << 16) / d->width;
            K.y_diff = (d->luma_height << 16) / d->height;
        }
        K.rows0 = (K.height_src + 1) / 2;
        K.cpr = (K.width_src + 14) / 8;
        long long total = (long long) K.cpr * K.height_src * K.batch;
        if (total > 0) {
            MC_BADARG(K.height_src > 65535 || K.batch > 65535, "more than 65535 rows / frames in one launch");
            dim3 grd((K.cpr + 127) / 128, K.height_src, K.batch);
            if (dissolve)
                k_luma_yuv422<true><<<grd, 128, 0, ctx->stream>>>(K);
            else
                k_luma_yuv422<false><<<grd, 128, 0, ctx->stream>>>(K);
            rc = mc_after_launch(ctx, "k_luma_yuv422");
        }
    } else {
        // :496-499 / :620-623: sizes must match
        if (d->width_src != d->width_dest || d->height_src != d->height_dest || d->width_src == 0 || d->height_src == 0) {
            if (flag)
                mltcuda_free(ctx, flag);
            mltcuda_set_error("Unable to dissolve incompatible images");
            return MLTCUDA_E_BADARG;
        }
        K.width_src = d->width_dest, K.height_src = d->height_dest, K.width_dest = d->width_dest;
        if (!dissolve) {
            MC_BADARG(d->luma_width <= 0 || d->luma_height <= 0, "bad luma map size");
            K.x_diff = (d->luma_width << 16) / d->width_dest;
            K.y_diff = (d->luma_height << 16) / d->height_dest;
        }
        K.rows0 = (K.height_src + 1) / 2;
        const int ppt = fmt == MLTCUDA_FMT_RGBA ? 4 : 2;
        long long per_frame = ((long long) K.width_src * K.height_src + ppt - 1) / ppt;
        MC_BADARG(per_frame * ppt >= (1ll << 31) || K.batch > 65535, "frame too large / too many frames for one launch");
        dim3 blocks(mc_blocks(per_frame, 256), 1, K.batch), threads(256, 1, 1);
        K.cpr = 0;
        if (K.width_src % ppt == 0 && (K.height_src + 7) / 8 <= 65535) {
            K.cpr = K.width_src / ppt;
            blocks = dim3((K.cpr + 31) / 32, (K.height_src + 7) / 8, K.batch), threads = dim3(32, 8, 1);
        }
        const int sel = (fmt == MLTCUDA_FMT_RGBA ? 0 : 4) | (dissolve ? 2 : 0) | (K.cpr > 0 ? 1 : 0);
        switch (sel) {
        case 0: k_luma_rgba<uint8_t, false, false><<<blocks, threads, 0, ctx->stream>>>(K); break;
        case 1: k_luma_rgba<uint8_t, false, true><<<blocks, threads, 0, ctx->stream>>>(K); break;
        case 2: k_luma_rgba<uint8_t, true, false><<<blocks, threads, 0, ctx->stream>>>(K); break;
        case 3: k_luma_rgba<uint8_t, true, true><<<blocks, threads, 0, ctx->stream>>>(K); break;
        case 4: k_luma_rgba<uint16_t, false, false><<<blocks, threads, 0, ctx->stream>>>(K); break;
        case 5: k_luma_rgba<uint16_t, false, true><<<blocks, threads, 0, ctx->stream>>>(K); break;
        case 6: k_luma_rgba<uint16_t, true, false><<<blocks, threads, 0, ctx->stream>>>(K); break;
        default: k_luma_rgba<uint16_t, true, true><<<blocks, threads, 0, ctx->stream>>>(K); break;
        }
        rc = mc_after_launch(ctx, "k_luma_rgba");
    }
    if (flag)
        mltcuda_free(ctx, flag);
    return rc;
}
