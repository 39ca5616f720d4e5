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
#include "melf_common.cuh"
#include "melf_args.h"


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

}  // namespace

__global__ void __launch_bounds__(256)
melf_convert_vec_kernel(const MelfConvertArgs A)
{
    const int groups_per_row = A.w >> 4;
    const long long total = (long long)A.n * A.h * groups_per_row;
    for (long long gidx = (long long)blockIdx.x * blockDim.x + threadIdx.x; gidx < total;
         gidx += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(gidx % groups_per_row);
        const long long rowi = gidx / groups_per_row;
        const int y = (int)(rowi % A.h);
        const int f = (int)(rowi / A.h);
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
        for (int i = 0; i < 12; ++i) out[i] = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) { lw[q] = 0; mw[q] = 0; }
        const int x0 = g << 4;
#pragma unroll
        for (int p = 0; p < 16; ++p) {
            // bytes 3p, 3p+1, 3p+2 of the 48-byte group
            const int b0 = 3 * p, b1 = 3 * p + 1, b2 = 3 * p + 2;
            const int B = (in[b0 >> 2] >> (8 * (b0 & 3))) & 255;
            const int G = (in[b1 >> 2] >> (8 * (b1 & 3))) & 255;
            const int R = (in[b2 >> 2] >> (8 * (b2 & 3))) & 255;
            const bool tail = A.s_tail && melf_is_tail_col(x0 + p, A.w);
            const uint32_t c = melf_hls_pixel(B, G, R, tail, A.hue_shift);
            out[b0 >> 2] |= (c & 255) << (8 * (b0 & 3));
            out[b1 >> 2] |= ((c >> 8) & 255) << (8 * (b1 & 3));
            out[b2 >> 2] |= ((c >> 16) & 255) << (8 * (b2 & 3));
            lw[p >> 2] |= ((c >> 8) & 255) << (8 * (p & 3));
            mw[p >> 2] |= in_range255(c, A.lo, A.hi) << (8 * (p & 3));
        }
        const size_t px = ((size_t)f * A.h + y) * A.w + x0;
        if (A.hls) {
            uint4 *dst = reinterpret_cast<uint4 *>(A.hls + px * 3);
            stg_stream(dst, make_uint4(out[0], out[1], out[2], out[3]));
            stg_stream(dst + 1, make_uint4(out[4], out[5], out[6], out[7]));
            stg_stream(dst + 2, make_uint4(out[8], out[9], out[10], out[11]));
        }
        if (A.light) stg_stream(reinterpret_cast<uint4 *>(A.light + px), make_uint4(lw[0], lw[1], lw[2], lw[3]));
        if (A.mask) stg_stream(reinterpret_cast<uint4 *>(A.mask + px), make_uint4(mw[0], mw[1], mw[2], mw[3]));
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

cudaError_t melf_launch_convert(const MelfConvertArgs &a, int num_sms, cudaStream_t stream)
{
    auto aligned16 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool vec = (a.w % 16 == 0) && aligned16(a.bgr) && (a.row_stride % 16 == 0) &&
                     (a.frame_stride % 16 == 0) && aligned16(a.hls) && aligned16(a.light) &&
                     aligned16(a.mask);
    if (vec) {
        const long long total = (long long)a.n * a.h * (a.w / 16);
        long long blocks = (total + 255) / 256;
        const long long cap = (long long)num_sms * 8;         // 8 resident CTAs of 256 per SM
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        melf_convert_vec_kernel<<<(unsigned)blocks, 256, 0, stream>>>(a);
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
