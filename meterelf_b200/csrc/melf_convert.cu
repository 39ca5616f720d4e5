// Fused colour pass over whole frames: BGR -> HLS_FULL + hue shift (+ lightness plane)
// (+ colour-range mask) in ONE sweep.
//
// Replaces convert_to_hls (_utils.py:100-102: cv2.cvtColor + numpy u8 add),
// cv2.split(...)[1] (_image.py:59) and get_mask_by_color (_utils.py:113-119: cv2.inRange)
// which the reference runs as four separate passes over the image.
//
// HBM-bound: 3 B/px in, up to 3 + 1 + 1 B/px out.  Vector path: each thread owns 16
// consecutive pixels = three 128-bit loads, and writes three 128-bit HLS stores, one
// 128-bit lightness store and one 128-bit mask store; consecutive threads own
// consecutive 48-byte groups so every warp request is fully coalesced.  The scalar path
// (one pixel per thread) covers unaligned rows / widths that are not a multiple of 16.
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "melf_common.cuh"
#include "melf_args.h"


#ifndef MELF_CONVERT_MIN_BLOCKS
#define MELF_CONVERT_MIN_BLOCKS 4      // occupancy 4..6 measured within 1 % of each other (issue-bound)
#endif

namespace {

__device__ __forceinline__ uint32_t in_range255(uint32_t c, uint32_t lo, uint32_t hi)
{
    const uint32_t H = c & 255, L = (c >> 8) & 255, S = (c >> 16) & 255;
    const bool ok = H >= (lo & 255) && H <= (hi & 255) && L >= ((lo >> 8) & 255) && L <= ((hi >> 8) & 255) &&
                    S >= ((lo >> 16) & 255) && S <= ((hi >> 16) & 255);
    return ok ? 255u : 0u;
}

__device__ __forceinline__ uint4 ldg_stream(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(uint4 *p, const uint4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// bytes 0 / 1 = the sign bit of a / b replicated over the byte (PRMT sign mode: selector
// nibble 8 + k); written as PTX because __byte_perm documents only 3 selector bits
__device__ __forceinline__ uint32_t prmt_sign(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, 0x00FB;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}

}  // namespace

// Vector kernel: 16 pixels (three 128-bit loads) per thread per iteration, grid-stride.
// Width is a multiple of 16 here, so no OpenCV scalar-tail column exists (tails are the
// last w mod 8 columns of a 256-pixel block) and melf_hls_fast applies.
// Mask test in the float pipe: the channel values are integer-valued floats M + v
// (M = 1.5 * 2^23), so a = v - lo and b = v - hi are exact packed subtractions and
// e = a * b - 0.5 is exact, never zero, and negative exactly for lo <= v <= hi.  The three
// sign bits are ANDed by one LOP3 and turned into the mask byte by PRMT's sign-replicate
// mode.  The hue shift is folded into the H bounds (a <= H <= b modulo 256); when that
// interval wraps, the complementary interval is tested with the sign reversed
// (e = -a * b + 0.5).
template <bool HLS, bool LIGHT, bool MASK>
__global__ void __launch_bounds__(256, MELF_CONVERT_MIN_BLOCKS)
melf_convert_vec_kernel(const MelfConvertArgs A)
{
    const uint32_t groups_per_row = (uint32_t)A.w >> 4;
    const uint32_t total = (uint32_t)A.n * (uint32_t)A.h * groups_per_row;     // < 2^32 (launcher)
    const float M = MELF_MAGIC_RINT;
    const int l0 = (A.lo >> 8) & 255, l1 = (A.hi >> 8) & 255, s0 = (A.lo >> 16) & 255, s1 = (A.hi >> 16) & 255;
    // H' = (H + shift) mod 256 in [lo, hi]  <=>  H in [a, b] mod 256
    const int ha = (int)((A.lo & 255) - (uint32_t)A.hue_shift) & 255;
    const int hb = (int)((A.hi & 255) - (uint32_t)A.hue_shift) & 255;
    const bool wrap = ha > hb;
    const bool full = wrap && ha == hb + 1;                           // complement empty: every H passes
    const int h0 = wrap ? hb + 1 : ha, h1 = wrap ? ha - 1 : hb;
    const float2 neg1 = melf_f2(-1.f);
    const float2 la_c = melf_f2(M + (float)l0), lb_c = melf_f2(M + (float)l1);
    const float2 sa_c = melf_f2(M + (float)s0), sb_c = melf_f2(M + (float)s1);
    // a_h = +-(V - (M + h0)) as one fma: V * hs - hs * (M + h0)
    const float hs = (wrap && !full) ? -1.f : 1.f;
    const float2 hs2 = melf_f2(hs), ha_c = melf_f2(-hs * (M + (float)h0)), hb_c = melf_f2(M + (float)h1);
    const float2 half_neg = melf_f2(-0.5f);
    const float2 eh_c = melf_f2(full ? -1e9f : (wrap ? 0.5f : -0.5f));
    for (uint32_t gidx = blockIdx.x * blockDim.x + threadIdx.x; gidx < total;
         gidx += gridDim.x * blockDim.x) {
        const uint32_t rowi = melf_fastdiv(gidx, A.gdiv);
        const uint32_t g = gidx - rowi * groups_per_row;
        const uint32_t f = melf_fastdiv(rowi, A.hdiv);
        const uint32_t y = rowi - f * (uint32_t)A.h;
        const uint4 *src = reinterpret_cast<const uint4 *>(
            A.bgr + (size_t)f * A.frame_stride + (size_t)y * A.row_stride) + 3 * g;
        uint32_t in[12];
        {
            const uint4 v0 = ldg_stream(src), v1 = ldg_stream(src + 1), v2 = ldg_stream(src + 2);
            in[0] = v0.x; in[1] = v0.y; in[2] = v0.z; in[3] = v0.w;
            in[4] = v1.x; in[5] = v1.y; in[6] = v1.z; in[7] = v1.w;
            in[8] = v2.x; in[9] = v2.y; in[10] = v2.z; in[11] = v2.w;
        }
        uint32_t out[12], lw[4], mw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t c[4];                       // per pixel: H | L << 8 | S << 16
            uint32_t lb[4], ob[4];
#pragma unroll
            for (int pr = 0; pr < 2; ++pr) {     // two pixels per pass through the packed pipe
                const int b0 = 3 * (4 * q + 2 * pr), b1 = b0 + 3;
                float2 Hf, Lf, Sf;
                melf_hls_fast2(
                    make_float2(melf_byte_biased(in[b0 >> 2], b0 & 3), melf_byte_biased(in[b1 >> 2], b1 & 3)),
                    make_float2(melf_byte_biased(in[(b0 + 1) >> 2], (b0 + 1) & 3),
                                melf_byte_biased(in[(b1 + 1) >> 2], (b1 + 1) & 3)),
                    make_float2(melf_byte_biased(in[(b0 + 2) >> 2], (b0 + 2) & 3),
                                melf_byte_biased(in[(b1 + 2) >> 2], (b1 + 2) & 3)), Hf, Lf, Sf);
                const int i = 2 * pr;
                lb[i] = __float_as_uint(Lf.x); lb[i + 1] = __float_as_uint(Lf.y);
                if (HLS) {
                    const uint32_t h0 = __float_as_uint(Hf.x) + (uint32_t)A.hue_shift,
                                   h1 = __float_as_uint(Hf.y) + (uint32_t)A.hue_shift;
                    c[i] = __byte_perm(__byte_perm(h0, lb[i], 0x0040), __float_as_uint(Sf.x), 0x0410);
                    c[i + 1] = __byte_perm(__byte_perm(h1, lb[i + 1], 0x0040), __float_as_uint(Sf.y), 0x0410);
                }
                if (MASK) {
                    const float2 eh = melf_fma2(melf_fma2(Hf, hs2, ha_c), melf_fma2(hb_c, neg1, Hf), eh_c);
                    const float2 el = melf_fma2(melf_fma2(la_c, neg1, Lf), melf_fma2(lb_c, neg1, Lf), half_neg);
                    const float2 es = melf_fma2(melf_fma2(sa_c, neg1, Sf), melf_fma2(sb_c, neg1, Sf), half_neg);
                    ob[i] = __float_as_uint(eh.x) & __float_as_uint(el.x) & __float_as_uint(es.x);
                    ob[i + 1] = __float_as_uint(eh.y) & __float_as_uint(el.y) & __float_as_uint(es.y);
                }
            }
            if (MASK)       // byte i = 0xFF where the sign of ob[i] is set
                mw[q] = __byte_perm(prmt_sign(ob[0], ob[1]), prmt_sign(ob[2], ob[3]), 0x5410);
            if (HLS) {
                out[3 * q] = __byte_perm(c[0], c[1], 0x4210);
                out[3 * q + 1] = __byte_perm(c[1], c[2], 0x5421);
                out[3 * q + 2] = __byte_perm(c[2], c[3], 0x6542);
            }
            if (LIGHT)
                lw[q] = __byte_perm(__byte_perm(lb[0], lb[1], 0x0040), __byte_perm(lb[2], lb[3], 0x0040), 0x5410);
        }
        const size_t px = (size_t)gidx << 4;        // == ((f * h + y) * w + 16 g): outputs are dense
        if (HLS) {
            uint4 *dst = reinterpret_cast<uint4 *>(A.hls + px * 3);
            stg_stream(dst, make_uint4(out[0], out[1], out[2], out[3]));
            stg_stream(dst + 1, make_uint4(out[4], out[5], out[6], out[7]));
            stg_stream(dst + 2, make_uint4(out[8], out[9], out[10], out[11]));
        }
        if (LIGHT) stg_stream(reinterpret_cast<uint4 *>(A.light + px), make_uint4(lw[0], lw[1], lw[2], lw[3]));
        if (MASK) stg_stream(reinterpret_cast<uint4 *>(A.mask + px), make_uint4(mw[0], mw[1], mw[2], mw[3]));
    }
}


// ---------------------------------------------------------------------------------------
// Table kernel (the large-batch path).  ncu on the arithmetic kernel above
// (profiles/r01_v3_ncu_convert_fma.txt) shows it bound by instruction issue, the half-rate ALU
// pipe and the packed-FMA pipe both ~58 % busy, DRAM at 35 %.  Two observations cut the
// arithmetic by more than half without changing a single output byte:
//  * L and S depend only on (max, min) of the three channel bytes: 2^16 pairs.  A static
//    table built once per device with the exact cv2-order scalar code (melf_hls_pixel) holds
//    them; each CTA stages it in shared memory as int16 entries L | inrange(L, S) << 15, so
//    one sign-extending 16-bit shared load yields the lightness byte AND the L/S part of the
//    colour-range test (the sign bit, ANDed with the hue test's sign bit).
//  * H = rint(256/360 * hue) is the round-to-nearest of the exact rational
//    256 * N / (3 * D), D = max - min <= 255.  Its fractional part is a multiple of 1/(3 D)
//    and can never be 1/2 (that would need 2^9 | 3 D (2m + 1)), so it stays at least
//    1/(6 D) >= 6.5e-4 away from a rounding tie, while cv2's float32 evaluation is within
//    ~1.3e-3 / D of the exact value and the float32 evaluation of
//    t * (1 / D) * (128 / 3) + sector offset used here within ~1e-4.  Hence the short form below gives cv2's byte for every colour; this is
//    CHECKED, not assumed: test_convert_hls_full_colour_cube_bit_exact runs all 2^24 colours
//    through this kernel against the oracle.
// One CTA of 1024 threads per SM (128 KB table, + 64 KB S table when HLS is written).
// ---------------------------------------------------------------------------------------
constexpr int TAB_THREADS = 1024;
constexpr int TAB_SMEM_LS = 65536 * 2;
constexpr int TAB_SMEM_S = 65536;

__global__ void melf_convert_build_table_kernel(uint16_t *ls)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // imax << 8 | imin
    const int imax = idx >> 8, imin = idx & 255;
    uint32_t v = 0;
    if (imin <= imax) {
        const uint32_t c = melf_hls_pixel(imin, imin, imax, false, 0);
        v = ((c >> 8) & 255u) | (((c >> 16) & 255u) << 8);
    }
    ls[idx] = (uint16_t)v;
}

extern __shared__ __align__(16) uint8_t melf_conv_smem[];

template <bool HLS, bool LIGHT, bool MASK>
__global__ void __launch_bounds__(TAB_THREADS, 1)
melf_convert_tab_kernel(const MelfConvertArgs A, const uint16_t *__restrict__ ls_table)
{
    int16_t *tab = reinterpret_cast<int16_t *>(melf_conv_smem);
    uint8_t *tab_s = melf_conv_smem + TAB_SMEM_LS;
    const int l0 = (A.lo >> 8) & 255, l1 = (A.hi >> 8) & 255, s0 = (A.lo >> 16) & 255, s1 = (A.hi >> 16) & 255;
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(ls_table);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tab);
        for (int i = threadIdx.x; i < 32768; i += TAB_THREADS) {
            const uint32_t two = __ldg(src + i);
            uint32_t out = 0;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int L = (two >> (16 * k)) & 255, S = (two >> (16 * k + 8)) & 255;
                const bool in = L >= l0 && L <= l1 && S >= s0 && S <= s1;
                out |= ((uint32_t)L | (in ? 0x8000u : 0u)) << (16 * k);
                if (HLS) tab_s[2 * i + k] = (uint8_t)S;
            }
            dst[i] = out;
        }
    }
    __syncthreads();

    const uint32_t groups_per_row = (uint32_t)A.w >> 4;
    const uint32_t total = (uint32_t)A.n * (uint32_t)A.h * groups_per_row;     // < 2^32 (launcher)
    const float M = MELF_MAGIC_RINT;
    // H' = (H + shift) mod 256 in [lo, hi]  <=>  H in [a, b] mod 256 (see the kernel above)
    const int ha = (int)((A.lo & 255) - (uint32_t)A.hue_shift) & 255;
    const int hb = (int)((A.hi & 255) - (uint32_t)A.hue_shift) & 255;
    const bool wrap = ha > hb;
    const bool full = wrap && ha == hb + 1;
    const int h0 = wrap ? hb + 1 : ha, h1 = wrap ? ha - 1 : hb;
    const float2 neg1 = melf_f2(-1.f);
    const float hs = (wrap && !full) ? -1.f : 1.f;
    const float2 hs2 = melf_f2(hs), ha_c = melf_f2(-hs * (M + (float)h0)), hb_c = melf_f2(M + (float)h1);
    const float2 eh_c = melf_f2(full ? -1e9f : (wrap ? 0.5f : -0.5f));
    const float2 tiny = melf_f2(1e-30f), c43 = melf_f2(128.f / 3.f), M2 = melf_f2(M);
    for (uint32_t gidx = blockIdx.x * blockDim.x + threadIdx.x; gidx < total;
         gidx += gridDim.x * blockDim.x) {
        const uint32_t rowi = melf_fastdiv(gidx, A.gdiv);
        const uint32_t g = gidx - rowi * groups_per_row;
        const uint32_t f = melf_fastdiv(rowi, A.hdiv);
        const uint32_t y = rowi - f * (uint32_t)A.h;
        const uint4 *src = reinterpret_cast<const uint4 *>(
            A.bgr + (size_t)f * A.frame_stride + (size_t)y * A.row_stride) + 3 * g;
        uint32_t in[12];
        {
            const uint4 v0 = ldg_stream(src), v1 = ldg_stream(src + 1), v2 = ldg_stream(src + 2);
            in[0] = v0.x; in[1] = v0.y; in[2] = v0.z; in[3] = v0.w;
            in[4] = v1.x; in[5] = v1.y; in[6] = v1.z; in[7] = v1.w;
            in[8] = v2.x; in[9] = v2.y; in[10] = v2.z; in[11] = v2.w;
        }
        uint32_t out[12], lw[4], mw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t c[4], lb[4], ob[4];
#pragma unroll
            for (int pr = 0; pr < 2; ++pr) {
                const int b0 = 3 * (4 * q + 2 * pr), b1 = b0 + 3;
                const float2 fb = make_float2(melf_byte_biased(in[b0 >> 2], b0 & 3), melf_byte_biased(in[b1 >> 2], b1 & 3));
                const float2 fg = make_float2(melf_byte_biased(in[(b0 + 1) >> 2], (b0 + 1) & 3),
                                              melf_byte_biased(in[(b1 + 1) >> 2], (b1 + 1) & 3));
                const float2 fr = make_float2(melf_byte_biased(in[(b0 + 2) >> 2], (b0 + 2) & 3),
                                              melf_byte_biased(in[(b1 + 2) >> 2], (b1 + 2) & 3));
                const float2 vmax = make_float2(fmaxf(fmaxf(fb.x, fg.x), fr.x), fmaxf(fmaxf(fb.y, fg.y), fr.y));
                const float2 vmin = make_float2(fminf(fminf(fb.x, fg.x), fr.x), fminf(fminf(fb.y, fg.y), fr.y));
                // imax << 8 | imin from the low mantissa bytes (bytes 1, 2 of the biased float are 0)
                const uint32_t i0 = __byte_perm(__float_as_uint(vmin.x), __float_as_uint(vmax.x), 0x1140);
                const uint32_t i1 = __byte_perm(__float_as_uint(vmin.y), __float_as_uint(vmax.y), 0x1140);
                const int e0 = tab[i0], e1 = tab[i1];                 // sign-extending 16-bit loads
                const float2 dl = melf_add2(melf_sub2(vmax, vmin), tiny);
                float2 inv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv.x) : "f"(dl.x));
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv.y) : "f"(dl.y));
                const float2 u_r = melf_sub2(fg, fb), u_g = melf_sub2(fb, fr), u_b = melf_sub2(fr, fg);
                const bool rx = vmax.x == fr.x, gx = vmax.x == fg.x, ry = vmax.y == fr.y, gy = vmax.y == fg.y;
                const float2 t = make_float2(rx ? u_r.x : gx ? u_g.x : u_b.x, ry ? u_r.y : gy ? u_g.y : u_b.y);
                const float2 off = make_float2(rx ? (u_r.x < 0.f ? 256.f : 0.f) : gx ? 256.f / 3.f : 512.f / 3.f,
                                               ry ? (u_r.y < 0.f ? 256.f : 0.f) : gy ? 256.f / 3.f : 512.f / 3.f);
                const float2 v = melf_fma2(melf_mul2(t, inv), c43, off);
                const float2 Hf = melf_add2(make_float2(fminf(v.x, 255.f), fminf(v.y, 255.f)), M2);
                const int i = 2 * pr;
                lb[i] = (uint32_t)e0; lb[i + 1] = (uint32_t)e1;
                if (HLS) {
                    const uint32_t hh0 = __float_as_uint(Hf.x) + (uint32_t)A.hue_shift,
                                   hh1 = __float_as_uint(Hf.y) + (uint32_t)A.hue_shift;
                    c[i] = __byte_perm(__byte_perm(hh0, lb[i], 0x0040), (uint32_t)tab_s[i0], 0x0410);
                    c[i + 1] = __byte_perm(__byte_perm(hh1, lb[i + 1], 0x0040), (uint32_t)tab_s[i1], 0x0410);
                }
                if (MASK) {
                    const float2 eh = melf_fma2(melf_fma2(Hf, hs2, ha_c), melf_fma2(hb_c, neg1, Hf), eh_c);
                    ob[i] = __float_as_uint(eh.x) & (uint32_t)e0;
                    ob[i + 1] = __float_as_uint(eh.y) & (uint32_t)e1;
                }
            }
            if (MASK) mw[q] = __byte_perm(prmt_sign(ob[0], ob[1]), prmt_sign(ob[2], ob[3]), 0x5410);
            if (HLS) {
                out[3 * q] = __byte_perm(c[0], c[1], 0x4210);
                out[3 * q + 1] = __byte_perm(c[1], c[2], 0x5421);
                out[3 * q + 2] = __byte_perm(c[2], c[3], 0x6542);
            }
            if (LIGHT)
                lw[q] = __byte_perm(__byte_perm(lb[0], lb[1], 0x0040), __byte_perm(lb[2], lb[3], 0x0040), 0x5410);
        }
        const size_t px = (size_t)gidx << 4;
        if (HLS) {
            uint4 *dst = reinterpret_cast<uint4 *>(A.hls + px * 3);
            stg_stream(dst, make_uint4(out[0], out[1], out[2], out[3]));
            stg_stream(dst + 1, make_uint4(out[4], out[5], out[6], out[7]));
            stg_stream(dst + 2, make_uint4(out[8], out[9], out[10], out[11]));
        }
        if (LIGHT) stg_stream(reinterpret_cast<uint4 *>(A.light + px), make_uint4(lw[0], lw[1], lw[2], lw[3]));
        if (MASK) stg_stream(reinterpret_cast<uint4 *>(A.mask + px), make_uint4(mw[0], mw[1], mw[2], mw[3]));
    }
}

__global__ void __launch_bounds__(256)
melf_convert_scalar_kernel(const MelfConvertArgs A)
{
    const long long total = (long long)A.n * A.h * A.w;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % A.w);
        const long long rowi = i / A.w;
        const int y = (int)(rowi % A.h);
        const int f = (int)(rowi / A.h);
        const uint8_t *p = A.bgr + (size_t)f * A.frame_stride + (size_t)y * A.row_stride + 3 * (size_t)x;
        const uint32_t c = melf_hls_pixel(p[0], p[1], p[2], A.s_tail && melf_is_tail_col(x, A.w), A.hue_shift);
        if (A.hls) {
            uint8_t *q = A.hls + (size_t)i * 3;
            q[0] = c & 255; q[1] = (c >> 8) & 255; q[2] = (c >> 16) & 255;
        }
        if (A.light) A.light[i] = (c >> 8) & 255;
        if (A.mask) A.mask[i] = (uint8_t)in_range255(c, A.lo, A.hi);
    }
}

// the static (L, S) table of the current device, built on first use
static cudaError_t convert_table(const uint16_t **out, cudaStream_t stream)
{
    static std::mutex mutex;
    static uint16_t *tables[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mutex);
    if (!tables[dev]) {
        uint16_t *t = nullptr;
        e = cudaMalloc(&t, 65536 * sizeof(uint16_t));
        if (e != cudaSuccess) return e;
        melf_convert_build_table_kernel<<<256, 256, 0, stream>>>(t);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);    // once: other streams may use it next
        if (e != cudaSuccess) { cudaFree(t); return e; }
        tables[dev] = t;
    }
    *out = tables[dev];
    return cudaSuccess;
}

template <bool HLS, bool LIGHT, bool MASK>
static cudaError_t launch_tab_one(const MelfConvertArgs &b, const uint16_t *table, unsigned grid,
                                  cudaStream_t stream)
{
    const int smem = TAB_SMEM_LS + (HLS ? TAB_SMEM_S : 0);
    // per device and cheap: set on every launch rather than caching per process
    cudaError_t e = cudaFuncSetAttribute(melf_convert_tab_kernel<HLS, LIGHT, MASK>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    melf_convert_tab_kernel<HLS, LIGHT, MASK><<<grid, TAB_THREADS, smem, stream>>>(b, table);
    return cudaGetLastError();
}

static cudaError_t launch_tab(const MelfConvertArgs &b, const uint16_t *table, int num_sms,
                              long long total_groups, cudaStream_t stream)
{
    long long blocks = (total_groups + TAB_THREADS - 1) / TAB_THREADS;
    if (blocks > num_sms) blocks = num_sms;
    if (blocks < 1) blocks = 1;
    const unsigned grid = (unsigned)blocks;
    switch ((b.hls ? 4 : 0) | (b.light ? 2 : 0) | (b.mask ? 1 : 0)) {
    case 7: return launch_tab_one<true, true, true>(b, table, grid, stream);
    case 6: return launch_tab_one<true, true, false>(b, table, grid, stream);
    case 5: return launch_tab_one<true, false, true>(b, table, grid, stream);
    case 4: return launch_tab_one<true, false, false>(b, table, grid, stream);
    case 3: return launch_tab_one<false, true, true>(b, table, grid, stream);
    case 2: return launch_tab_one<false, true, false>(b, table, grid, stream);
    case 1: return launch_tab_one<false, false, true>(b, table, grid, stream);
    default: return cudaSuccess;
    }
}

static void make_fastdiv(uint32_t d, uint32_t (&out)[3])
{
    uint32_t l = 0;
    while ((1ull << l) < d) ++l;
    out[0] = (uint32_t)(((1ull << 32) * ((1ull << l) - d)) / d + 1);
    out[1] = l < 1 ? l : 1;
    out[2] = l > 0 ? l - 1 : 0;
}

cudaError_t melf_launch_convert(const MelfConvertArgs &a_in, int num_sms, cudaStream_t stream)
{
    MelfConvertArgs a = a_in;
    make_fastdiv((uint32_t)(a.w / 16 > 0 ? a.w / 16 : 1), a.gdiv);
    make_fastdiv((uint32_t)a.h, a.hdiv);
    // the vector kernel indexes 16-pixel groups with 32 bits: split very large batches
    const long long per_frame = (long long)a.h * (a.w / 16 > 0 ? a.w / 16 : 1);
    const long long max_frames = ((1ll << 31) / per_frame) > 0 ? ((1ll << 31) / per_frame) : 1;
    if (a.n > max_frames) {
        for (long long f0 = 0; f0 < a.n; f0 += max_frames) {
            MelfConvertArgs part = a_in;
            part.n = (int32_t)((a.n - f0) < max_frames ? (a.n - f0) : max_frames);
            part.bgr = a.bgr + (size_t)f0 * a.frame_stride;
            const size_t px0 = (size_t)f0 * a.h * a.w;
            if (part.hls) part.hls += px0 * 3;
            if (part.light) part.light += px0;
            if (part.mask) part.mask += px0;
            cudaError_t e = melf_launch_convert(part, num_sms, stream);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    }
    auto aligned16 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool vec = (a.w % 16 == 0) && aligned16(a.bgr) && (a.row_stride % 16 == 0) &&
                     (a.frame_stride % 16 == 0) && aligned16(a.hls) && aligned16(a.light) &&
                     aligned16(a.mask);
    if (vec) {
        const long long total = (long long)a.n * a.h * (a.w / 16);
        long long blocks = (total + 255) / 256;
        const long long cap = (long long)num_sms * 8;         // measured best: ~2 waves of resident CTAs, grid-stride
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        MelfConvertArgs b = a;
        // an empty range on any channel (lo > hi) selects nothing, like cv2.inRange
        const bool empty = (b.lo & 255) > (b.hi & 255) || ((b.lo >> 8) & 255) > ((b.hi >> 8) & 255) ||
                           ((b.lo >> 16) & 255) > ((b.hi >> 16) & 255);
        if (empty) { b.lo = 0x000001; b.hi = 0x000000; }
        const unsigned gb = (unsigned)blocks;
        const int sel = (b.hls ? 4 : 0) | (b.light ? 2 : 0) | (b.mask ? 1 : 0);
        if (empty && b.mask) {
            // handled by the kernel's unsigned compare only for non-empty ranges
            cudaError_t e = cudaMemsetAsync(b.mask, 0, (size_t)b.n * b.h * b.w, stream);
            if (e != cudaSuccess) return e;
            b.mask = nullptr;
        }
        // large batches: the table kernel (one persistent CTA per SM); small ones are not worth
        // staging 128 KB per CTA.  MELF_CONVERT_IMPL=fma|tab forces one of them.
        const char *impl = std::getenv("MELF_CONVERT_IMPL");
        const bool want_tab = impl ? std::strcmp(impl, "tab") == 0
                                   : (long long)b.n * b.h * b.w >= (long long)num_sms * TAB_THREADS * 64;
        if (want_tab && !(impl && std::strcmp(impl, "fma") == 0)) {
            const uint16_t *table = nullptr;
            cudaError_t e = convert_table(&table, stream);
            if (e != cudaSuccess) return e;
            return launch_tab(b, table, num_sms, total, stream);
        }
        switch ((b.hls ? 4 : 0) | (b.light ? 2 : 0) | (b.mask ? 1 : 0)) {
        case 7: melf_convert_vec_kernel<true, true, true><<<gb, 256, 0, stream>>>(b); break;
        case 6: melf_convert_vec_kernel<true, true, false><<<gb, 256, 0, stream>>>(b); break;
        case 5: melf_convert_vec_kernel<true, false, true><<<gb, 256, 0, stream>>>(b); break;
        case 4: melf_convert_vec_kernel<true, false, false><<<gb, 256, 0, stream>>>(b); break;
        case 3: melf_convert_vec_kernel<false, true, true><<<gb, 256, 0, stream>>>(b); break;
        case 2: melf_convert_vec_kernel<false, true, false><<<gb, 256, 0, stream>>>(b); break;
        case 1: melf_convert_vec_kernel<false, false, true><<<gb, 256, 0, stream>>>(b); break;
        default: break;
        }
        (void)sel;
    } else {
        const long long total = (long long)a.n * a.h * a.w;
        long long blocks = (total + 255) / 256;
        const long long cap = (long long)num_sms * 8;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        melf_convert_scalar_kernel<<<(unsigned)blocks, 256, 0, stream>>>(a);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// cv2.inRange on a 3-channel u8 image (get_mask_by_color on an HLS image, _utils.py:113-119):
// 3 B/px in, 1 B/px out, HBM-bound; 16 pixels per thread with 128-bit loads when the row is aligned
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
melf_in_range_kernel(const uint8_t *__restrict__ img, size_t row_stride, size_t frame_stride, int n, int h, int w,
                     uint32_t lo, uint32_t hi, uint8_t *__restrict__ mask)
{
    const int groups = (w + 15) >> 4;
    const long long total = (long long)n * h * groups;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(i % groups);
        const long long rowi = i / groups;
        const int y = (int)(rowi % h), f = (int)(rowi / h);
        const uint8_t *src = img + (size_t)f * frame_stride + (size_t)y * row_stride + 48 * (size_t)g;
        uint8_t *dst = mask + ((size_t)f * h + y) * (size_t)w + 16 * (size_t)g;
        const int npx = min(16, w - 16 * g);
        if (npx == 16 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
            const uint4 *v = reinterpret_cast<const uint4 *>(src);
            const uint4 a = __ldg(v), b = __ldg(v + 1), c = __ldg(v + 2);
            const uint32_t wv[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
            uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
            for (int px = 0; px < 16; ++px) {
                const int b0 = 3 * px;
                // bytes b0 .. b0+2 -> one packed H | L<<8 | S<<16 word
                const uint32_t lo_w = wv[b0 >> 2], hi_w = wv[(b0 + 2) >> 2];
                const uint32_t pix = __funnelshift_r(lo_w, hi_w, 8 * (b0 & 3)) & 0xffffffu;
                o[px >> 2] |= (in_range255(pix, lo, hi) & 255u) << (8 * (px & 3));
            }
            *reinterpret_cast<uint4 *>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
        } else {
            for (int px = 0; px < npx; ++px) {
                const uint32_t pix = src[3 * px] | (src[3 * px + 1] << 8) | (src[3 * px + 2] << 16);
                dst[px] = (uint8_t)in_range255(pix, lo, hi);
            }
        }
    }
}

cudaError_t melf_launch_in_range(const uint8_t *img, size_t row_stride, size_t frame_stride, int n, int h, int w,
                                 uint32_t lo, uint32_t hi, uint8_t *mask, int num_sms, cudaStream_t stream)
{
    const bool empty = (lo & 255) > (hi & 255) || ((lo >> 8) & 255) > ((hi >> 8) & 255) ||
                       ((lo >> 16) & 255) > ((hi >> 16) & 255);
    if (empty) return cudaMemsetAsync(mask, 0, (size_t)n * h * w, stream);     // like cv2.inRange
    const long long total = (long long)n * h * ((w + 15) >> 4);
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)num_sms * 8) blocks = (long long)num_sms * 8;
    if (blocks < 1) blocks = 1;
    melf_in_range_kernel<<<(unsigned)blocks, 256, 0, stream>>>(img, row_stride, frame_stride, n, h, w, lo, hi, mask);
    return cudaGetLastError();
}
