// Dial location on the 5th-generation tensor cores: TM_CCOEFF as an exact u8 x u8 -> s32
// implicit GEMM (tcgen05.mma.kind::i8, accumulators in TMEM), with the lightness conversion
// and window sums in a small preparation kernel in front of it and the score and arg-max
// fused behind it.
//
// Replaces, like melf_match.cu, ImageFile._find_dials (_image.py:57-66) = cv2.split(hls)[1]
// + match_template (_utils.py:91-97: cv2.matchTemplate(TM_CCOEFF) + cv2.minMaxLoc).
// ncu on the v1 dp4a kernel (profiles/r01_v1_ncu_full.txt) showed the correlation bound
// by the integer dot-product issue rate (fma pipe, DRAM at 0.5 %), i.e. contraction-bound,
// which is the condition BASELINE.json sets for moving it to the tensor pipe.
//
// Formulation.  For one template row i the correlation along x is a product with a
// Toeplitz matrix of that row:
//     R_i(y, x0+n) = sum_k L(y+i, x0+k) * Toep_i[k][n],   Toep_i[k][n] = T(i, k-n)
// (K = tw + 63 -> 256 for 64 result columns).  Two template rows share one MMA of N = 128:
//     D_q[r][n] += L(r + 2p, .) . Toep_{2p+q}[.][n]      q = 0, 1;  p = 0 .. ceil(th/2)-1
//     sum T*L (y, x) = D_0[y][x] + D_1[y+1][x]
// The A operand of step p is the SAME shared-memory image shifted down by 2p rows: the
// lightness slab is stored as 16-pixel-wide column strips ([strip][row][16 B]), which is
// the canonical K-major no-swizzle UMMA layout with an 8-row-group pitch of exactly 128 B,
// so a shift by any number of rows is a 16 B-granular change of the descriptor's start
// address.  The Toeplitz operands (119 rows -> 60 pairs x 32 KB) are built once per
// params on the host, live in L2 and stream through a 2-stage cp.async.bulk pipeline.
//
// A work item = one 64-column x <=137-row block of one frame's result map; for the sample
// geometry, 63 x 132, that is one whole frame.  Two kernels per batch:
//   melf_match_prep_kernel  one CTA per item: BGR crop -> lightness (bit-exact float32 order of
//       cv2, no XU pipe ops) -> slab in the strip layout + window sums, to global scratch
//       (64 KB + 34 KB per item; L2/HBM traffic that is small next to the MMA time);
//   melf_match_tc_kernel    persistent, one CTA per SM, two items per pass: slabs by one bulk
//       copy each, Toeplitz stream, 60 x (2 items x 2 M-tiles x 8 K-steps) tcgen05.mma of
//       128 x 128 x 32, then TMEM -> registers, exact score (N*sumTL - sumT*sumL)/N -> float32,
//       64-bit (score, ~raster index) key, warp/block max, one atomicMax per item.  The
//       operands of the next pass are fetched under the epilogue of the current one.
// Everything is integer-exact, so results are bit-identical to melf_match.cu and to the
// oracle.
#include <cstring>
#include <cuda.h>

#include "melf_common.cuh"
#include "melf_args.h"

#ifndef MELF_TC_TMEM_A
#define MELF_TC_TMEM_A 0    // 1: Toeplitz pair via tcgen05.cp into tensor memory -- measured slower (the copy serialises with the MMAs: 95 k vs 82 k clk per pass)
#endif

namespace {

#ifndef MELF_TC_THREADS
#define MELF_TC_THREADS 512
#endif
constexpr int TC_THREADS = MELF_TC_THREADS;
constexpr int SLAB_ROWS = 264;                       // rows of a lightness slab (118 + 144 -> 264)
constexpr int SLAB_DATA_ROWS = 256;                  // rows that can hold image data (plan cap)
constexpr int KS = 16;                               // 16-byte K strips: K = 256
constexpr int STRIP_BYTES = SLAB_ROWS * 16;          // 4224
constexpr int SLAB_BYTES = KS * STRIP_BYTES;         // 67584
constexpr int BM = 128;                              // MMA M: 64 result columns x two template rows
constexpr int BSTRIP_BYTES = BM * 16;                // 2048
constexpr int BSTAGE_BYTES = KS * BSTRIP_BYTES;      // 32768
constexpr int NSTAGE = 2;
constexpr int SROWS = 137;                           // max result rows of a work item
constexpr int NCOLS = 144;                           // TMEM columns of one accumulator = max MMA N
constexpr int TMEM_A_COL = 2 * NCOLS;                // Toeplitz pair in tensor memory: 2 buffers of
constexpr int TMEM_A_COLS = 64;                      //   128 lanes x 64 columns (= 128 x 256 bytes)
constexpr int WS_PITCH = 144;                        // window-sum row pitch (words)
constexpr int SMEM_SLABS = 0;
constexpr int SMEM_B = 2 * SLAB_BYTES;               // 135168
constexpr int SMEM_CTRL = SMEM_B + NSTAGE * BSTAGE_BYTES;   // 200704

// tcgen05 instruction descriptor: D = s32, A = B = u8, both K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t make_idesc(uint32_t n)
{
    return (2u << 4) | (0u << 7) | (0u << 10) | ((n >> 3) << 17) | ((128u >> 4) << 24);
}

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 :: "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// K-major, no swizzle: 8-row groups 128 B apart (SBO), the two 16-byte K chunks of one
// K = 32 instruction one strip apart (LBO); descriptor version 1 (sm_100)
__device__ __forceinline__ u64 umma_desc(uint32_t addr, uint32_t lbo_bytes)
{
    return (u64)((addr >> 4) & 0x3fffu) | ((u64)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((u64)((128u >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, u64 adesc, u64 bdesc, uint32_t idesc,
                                        uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}" :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// A operand from tensor memory (lane = row m, 8 columns = the 32 K bytes of one instruction)
__device__ __forceinline__ void umma_i8_ta(uint32_t tmem_d, uint32_t tmem_a, u64 bdesc, uint32_t idesc,
                                           uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t"
        "}" :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// shared -> tensor memory, 128 rows x 32 bytes (same canonical layout / descriptor as an MMA operand)
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t taddr, u64 sdesc)
{
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" :: "r"(taddr), "l"(sdesc) : "memory");
}

// one lane of a converged warp; with a warp-uniform branch around it the compiler keeps the
// descriptors in uniform registers and issues UTCIMMA back to back (a `tid == 32` branch makes it
// wrap every MMA in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop: ~94 clk per instruction issued,
// profiles/r01_v6_mma_rate.txt)
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[33])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld1(uint32_t taddr, uint32_t &v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// geometry of one work item
struct Item {
    int frame, x0, y0, w_c, h_c, n_cols;
    bool valid;
};

__device__ __forceinline__ Item get_item(const MelfMatchTcArgs &A, long long idx)
{
    Item it;
    it.valid = idx < A.n_items;
    if (!it.valid) { it.frame = 0; it.x0 = it.y0 = 0; it.w_c = it.h_c = 0; it.n_cols = 16; return it; }
    const int per_frame = A.n_tx * A.n_ty;
    it.frame = (int)(idx / per_frame);
    const int rem = (int)(idx - (long long)it.frame * per_frame);
    const int ty = rem / A.n_tx, tx = rem - ty * A.n_tx;
    it.x0 = tx * 64;
    it.y0 = ty * A.chunk_h;
    it.w_c = min(64, A.geom.res_w - it.x0);
    it.h_c = min(A.chunk_h, A.geom.res_h - it.y0);
    it.n_cols = (it.h_c + 1 + 15) & ~15;       // result rows 0..h_c-1 need D columns 0..h_c
    return it;
}

// 8388608 + v as float32 bits (0x4B0000vv) for byte `i` of `word`: int -> float without the
// conversion (XU) pipe; ordering of these floats = ordering of the bytes
__device__ __forceinline__ float byte_as_biased_float(uint32_t word, int i)
{
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7650u + (uint32_t)i));
}

// lightness of the pixel whose B, G, R bytes are bytes i0.. of the 12-word group `v`; same
// float32 operation order as melf_lightness() (= cv2's):
//     L = rint(((max/255 + min/255) * .5) * 255)
// (x * .5) * 255 and x * 127.5 are the same real number, hence the same rounding; the final
// round-half-even to integer is the rounding of the addition of 1.5 * 2^23.  Returns the
// float whose low mantissa byte is L.
__device__ __forceinline__ uint32_t lightness_bits(float fb, float fg, float fr)
{
    const float k = 1.f / 255.f;
    const float hi = fmaxf(fmaxf(fb, fg), fr), lo = fminf(fminf(fb, fg), fr);
    const float vmax = __fmul_rn(__fadd_rn(hi, -8388608.f), k);
    const float vmin = __fmul_rn(__fadd_rn(lo, -8388608.f), k);
    const float l255 = __fmul_rn(__fadd_rn(vmax, vmin), 127.5f);
    return __float_as_uint(__fadd_rn(l255, 12582912.f));
}

// 16 consecutive pixels of slab row r / strip s of an item -> one 16-byte shared store
__device__ __forceinline__ uint4 stage_unit(const MelfMatchTcArgs &A, const Item &it, int r, int s)
{
    uint4 out = make_uint4(0, 0, 0, 0);
    const int cy = it.y0 + r, cx = it.x0 + 16 * s;
    if (!it.valid || cy >= A.geom.crop_h || cx >= A.geom.crop_w) return out;
    const uint8_t *src = A.img + (size_t)it.frame * A.frame_stride +
                         (size_t)(A.geom.crop_y0 + cy) * A.row_stride +
                         (size_t)A.channels * (size_t)(A.geom.crop_x0 + cx);
    const uintptr_t addr = reinterpret_cast<uintptr_t>(src);
    const uint4 *al = reinterpret_cast<const uint4 *>(addr & ~(uintptr_t)15);
    const uint32_t off = (uint32_t)(addr & 15);
    const uint32_t ws = off >> 2, bs = (off & 3) * 8;
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
    const int nchunk = A.channels == 3 ? 4 : 2;
    uint32_t e[16];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c < nchunk && reinterpret_cast<uintptr_t>(al + c) < limit) v = __ldg(al + c);
        e[4 * c] = v.x; e[4 * c + 1] = v.y; e[4 * c + 2] = v.z; e[4 * c + 3] = v.w;
    }
    // shift right by `off` bytes: whole words by selection, the rest by funnel shift
    uint32_t t[13];
#pragma unroll
    for (int j = 0; j < 13; ++j)
        t[j] = ws == 0 ? e[j] : ws == 1 ? e[j + 1] : ws == 2 ? e[j + 2] : e[j + 3];
    uint32_t v[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(t[j], t[j + 1], bs);
    if (A.channels == 1) return make_uint4(v[0], v[1], v[2], v[3]);
    uint32_t o[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        uint32_t r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int b0 = 3 * (4 * g + q);            // byte index of B inside the 48-byte group
            r[q] = lightness_bits(byte_as_biased_float(v[b0 >> 2], b0 & 3),
                                  byte_as_biased_float(v[(b0 + 1) >> 2], (b0 + 1) & 3),
                                  byte_as_biased_float(v[(b0 + 2) >> 2], (b0 + 2) & 3));
        }
        const uint32_t lo = __byte_perm(r[0], r[1], 0x0040), hi = __byte_perm(r[2], r[3], 0x0040);
        o[g] = __byte_perm(lo, hi, 0x5410);
    }
    return make_uint4(o[0], o[1], o[2], o[3]);
}

}  // namespace


extern __shared__ __align__(1024) uint8_t melf_tc_smem[];

// ---------------------------------------------------------------------------------------
// Kernel 1 of 2: operand preparation.  One CTA per work item: BGR crop -> lightness slab in
// the UMMA strip layout (written to global scratch so the tensor-core kernel can fetch it with
// ONE 64 KB bulk copy) and the window sums S(y, x) = sum over the template footprint of L
// (the "mean(T) * sum_window L" term of TM_CCOEFF), x-major, to global scratch.
// HBM/L2-bound and small next to the correlation; keeping it out of the tensor-core kernel
// leaves that kernel's shared memory to the operands and its issue slots to the MMA stream
// (ncu on the fused v2 kernel: tensor pipe only 52 % active, profiles/r01_v3_ncu_match.txt).
// ---------------------------------------------------------------------------------------
constexpr int PREP_THREADS = 512;
constexpr int PREP_SMEM_H = SLAB_BYTES;                   // H table: 256 rows x 32 words (u16 pairs)
constexpr int PREP_SMEM_TOTAL = SLAB_BYTES + 32768;
constexpr int WSUM_WORDS = 64 * WS_PITCH;                 // per work item

__global__ void __launch_bounds__(PREP_THREADS, 2)
melf_match_prep_kernel(const MelfMatchTcArgs A)
{
    uint8_t *smem = melf_tc_smem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long li = blockIdx.x;                       // item index inside this launch
    const Item it = get_item(A, A.item0 + li);

    // ---- 1. lightness slab: warp-unit = 8 rows x 4 strips; lane = strip * 8 + row, so that
    // every quarter warp stores 128 contiguous bytes (conflict-free 128-bit shared stores)
    for (int u = warp; u < (SLAB_ROWS / 8) * 4; u += PREP_THREADS / 32) {
        const int rg = u >> 2, sg = u & 3;
        const int r = rg * 8 + (lane & 7), s = sg * 4 + (lane >> 3);
        *reinterpret_cast<uint4 *>(smem + s * STRIP_BYTES + r * 16) = stage_unit(A, it, r, s);
    }
    __syncthreads();

    // ---- 2. slab -> global scratch (coalesced) ------------------------------------------
    {
        uint4 *dst = reinterpret_cast<uint4 *>(A.slabs + (size_t)li * SLAB_BYTES);
        const uint4 *src = reinterpret_cast<const uint4 *>(smem);
        for (int i = tid; i < SLAB_BYTES / 16; i += PREP_THREADS) dst[i] = src[i];
    }

    // ---- 3a. horizontal window sums H(row, x) = sum_{j<tw} L(row, x+j), x < 64 ------------
    // thread = row; stored as u16 pairs, word (row, xp) at row*32 + (xp ^ (row & 31))
    if (tid < SLAB_DATA_ROWS) {
        const int row = tid;
        const uint8_t *sl = smem + row * 16;
        uint32_t *Hb = reinterpret_cast<uint32_t *>(smem + PREP_SMEM_H) + row * 32;
        uint32_t w[16];                       // bytes 0..63 of the row (the leaving bytes)
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const uint4 v = *reinterpret_cast<const uint4 *>(sl + s * STRIP_BYTES);
            w[4 * s] = v.x; w[4 * s + 1] = v.y; w[4 * s + 2] = v.z; w[4 * s + 3] = v.w;
        }
        int h0 = 0;
#pragma unroll
        for (int s = 0; s < KS; ++s) {
            const uint4 v = *reinterpret_cast<const uint4 *>(sl + s * STRIP_BYTES);
            h0 = (int)__dp4a(v.x, A.ones_words[4 * s], (uint32_t)h0);
            h0 = (int)__dp4a(v.y, A.ones_words[4 * s + 1], (uint32_t)h0);
            h0 = (int)__dp4a(v.z, A.ones_words[4 * s + 2], (uint32_t)h0);
            h0 = (int)__dp4a(v.w, A.ones_words[4 * s + 3], (uint32_t)h0);
        }
        // entering bytes: columns tw .. tw+62, i.e. 80 bytes from strip tw/16 on
        const int s0 = A.tmpl_w >> 4;
        uint32_t e[20];
#pragma unroll
        for (int s = 0; s < 5; ++s) {
            uint4 v = make_uint4(0, 0, 0, 0);
            if (s0 + s < KS) v = *reinterpret_cast<const uint4 *>(sl + (s0 + s) * STRIP_BYTES);
            e[4 * s] = v.x; e[4 * s + 1] = v.y; e[4 * s + 2] = v.z; e[4 * s + 3] = v.w;
        }
        const uint32_t o = (uint32_t)A.tmpl_w & 15, ws = o >> 2, bs = (o & 3) * 8;
        uint32_t t[17], g[16];
#pragma unroll
        for (int j = 0; j < 17; ++j)
            t[j] = ws == 0 ? e[j] : ws == 1 ? e[j + 1] : ws == 2 ? e[j + 2] : e[j + 3];
#pragma unroll
        for (int j = 0; j < 16; ++j) g[j] = __funnelshift_r(t[j], t[j + 1], bs);
        int h = h0;
        const int sw = row & 31;
#pragma unroll
        for (int xp = 0; xp < 32; ++xp) {
            const int xa = 2 * xp, xb = 2 * xp + 1;
            const uint32_t lo = (uint32_t)h;
            h += (int)((g[xa >> 2] >> (8 * (xa & 3))) & 255) - (int)((w[xa >> 2] >> (8 * (xa & 3))) & 255);
            const uint32_t hi = (uint32_t)h;
            h += (int)((g[xb >> 2] >> (8 * (xb & 3))) & 255) - (int)((w[xb >> 2] >> (8 * (xb & 3))) & 255);
            Hb[xp ^ sw] = lo | (hi << 16);
        }
    }
    __syncthreads();          // H complete; every read of the slab (copy-out, 3a) is done

    // ---- 3b. vertical: S(y, x) = sum_{i<th} H(y+i, x); thread = (segment, x pair) ---------
    // S goes into the (now dead) slab, x-major: S[x * WS_PITCH + y]
    {
        const int seg = tid >> 5, xp = lane;
        const uint32_t *Hb = reinterpret_cast<const uint32_t *>(smem + PREP_SMEM_H);
        uint32_t *S = reinterpret_cast<uint32_t *>(smem);
        const int rs = (it.h_c + 15) >> 4;
        const int ya = seg * rs, yb = min(it.h_c, ya + rs);
        if (ya < yb) {
            uint32_t s0 = 0, s1 = 0;
            for (int i = 0; i < A.tmpl_h; ++i) {
                const int row = ya + i;
                const uint32_t v = Hb[row * 32 + (xp ^ (row & 31))];
                s0 += v & 0xffff; s1 += v >> 16;
            }
            for (int y = ya; y < yb; ++y) {
                S[(2 * xp) * WS_PITCH + y] = s0;
                S[(2 * xp + 1) * WS_PITCH + y] = s1;
                const int rin = y + A.tmpl_h;
                const uint32_t vin = Hb[rin * 32 + (xp ^ (rin & 31))];      // row <= 255 always
                const uint32_t vout = Hb[y * 32 + (xp ^ (y & 31))];
                s0 += (vin & 0xffff) - (vout & 0xffff);
                s1 += (vin >> 16) - (vout >> 16);
            }
        }
    }
    __syncthreads();
    {
        uint32_t *dst = A.wsum + (size_t)li * WSUM_WORDS;
        const uint32_t *S = reinterpret_cast<const uint32_t *>(smem);
        for (int i = tid; i < WSUM_WORDS; i += PREP_THREADS) dst[i] = S[i];
    }
}

// ---------------------------------------------------------------------------------------
// Kernel 2 of 2: the correlation.  Persistent, one CTA per SM, two work items per pass
// (they share every Toeplitz stage, which halves the L2 -> SM operand traffic).
// The Toeplitz pair is the A operand (M = 128 = 64 result columns x 2 template rows, row
// m = 2 n + q), the lightness slab the B operand (N = result rows + 1, rounded up to 16), so
// the result-row extent costs exactly what it needs (132 rows -> N = 144) instead of two
// 128-row M tiles:
//     D[2n+q][r] += Toep_{2p+q}[n][.] . L(r + 2p, .)        p = 0 .. ceil(th/2)-1
//     sum T*L (y, x) = D[2x][y] + D[2x+1][y+1]
//   thread 0   producer: slabs (one 66 KB bulk copy each) and Toeplitz stages
//   thread 32  issuer:   n_pairs x (2 items x 8 K-steps) tcgen05.mma of 128 x N x 32
//   all warps  epilogue: window sums prefetched from global, TMEM -> registers (lane pair
//              2x / 2x+1 holds the two template-row halves of column x), exact integer
//              numerator, arg-max with float32 rounding only for close calls.
// As soon as the accumulators are complete the producer refills both slabs and the first
// Toeplitz stages for the NEXT pass, so those copies run under the epilogue.
// ---------------------------------------------------------------------------------------
constexpr int SMEM_TOTAL2 = SMEM_CTRL + 512;

struct Ctrl2 {
    unsigned long long full[NSTAGE];
    unsigned long long empty[NSTAGE];
    unsigned long long accum;
    unsigned long long slab_full;
    uint32_t tmem_base;
    uint32_t pad;
    u64 red[2][16];            // per pass parity
};

// exact TM_CCOEFF value of an integer numerator, rounded once to float32 (cv2's result type)
__device__ __forceinline__ float score_of(long long num, long long n_t)
{
    return __double2float_rn(__ddiv_rn((double)num, (double)n_t));
}

template <bool WRITE_MAP>
__global__ void __launch_bounds__(TC_THREADS, 1)
melf_match_tc_kernel(const MelfMatchTcArgs A)
{
    uint8_t *smem = melf_tc_smem;
    Ctrl2 *ctrl = reinterpret_cast<Ctrl2 *>(smem + SMEM_CTRL);
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);          // provably warp-uniform
    const uint32_t slab_addr = smem_u32(smem + SMEM_SLABS), b_addr = smem_u32(smem + SMEM_B);

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(smem_u32(&ctrl->full[s]), 1);
            mbar_init(smem_u32(&ctrl->empty[s]), 1);
        }
        mbar_init(smem_u32(&ctrl->accum), 1);
        mbar_init(smem_u32(&ctrl->slab_full), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     :: "r"(smem_u32(&ctrl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctrl->tmem_base;

    const long long n_local = A.n_items - A.item0 < A.max_items ? A.n_items - A.item0 : A.max_items;
    const long long n_pass = (n_local + 1) / 2;
    const int n_pre = A.n_pairs < NSTAGE ? A.n_pairs : NSTAGE;     // stages prefetched ahead of a pass

    // producer helpers (thread 0 only).  Stage g of the running stage counter uses ring slot
    // g % NSTAGE; its "empty" barrier has completed (g / NSTAGE) phases when it may be refilled.
    auto load_stage = [&](uint32_t g, int p) {
        const uint32_t st = g % NSTAGE, ph = (g / NSTAGE) & 1;
        mbar_wait(smem_u32(&ctrl->empty[st]), ph ^ 1);
        mbar_expect_tx(smem_u32(&ctrl->full[st]), BSTAGE_BYTES);
        bulk_load(b_addr + st * BSTAGE_BYTES, A.toeplitz + (size_t)p * BSTAGE_BYTES, BSTAGE_BYTES,
                  smem_u32(&ctrl->full[st]));
    };
    auto load_slabs = [&](long long pass) {
        const long long i0 = 2 * pass;
        const int n = (i0 + 1 < n_local) ? 2 : 1;
        mbar_expect_tx(smem_u32(&ctrl->slab_full), (uint32_t)(n * SLAB_BYTES));
        for (int k = 0; k < n; ++k)
            bulk_load(slab_addr + k * SLAB_BYTES, A.slabs + (size_t)(i0 + k) * SLAB_BYTES, SLAB_BYTES,
                      smem_u32(&ctrl->slab_full));
    };

    uint32_t stage_count = 0;       // Toeplitz stages consumed before this pass
    uint32_t pass_count = 0;
    if (tid == 0 && (long long)blockIdx.x < n_pass) {
        load_slabs(blockIdx.x);
        for (int p = 0; p < n_pre; ++p) load_stage((uint32_t)p, p);
    }

    for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x, ++pass_count) {
        Item items[2];
        items[0] = get_item(A, A.item0 + 2 * pass);
        items[1] = get_item(A, A.item0 + 2 * pass + 1);
        if (2 * pass + 1 >= n_local) items[1].valid = false;
        const bool has_next = pass + gridDim.x < n_pass;
        long long *tl = (A.timeline && blockIdx.x == 0 && pass_count < 8) ? A.timeline + pass_count * 8 : nullptr;

        // ---- the correlation: Toeplitz stream + MMA ------------------------------------------
        if (tid == 0) {
            for (int p = n_pre; p < A.n_pairs; ++p) load_stage(stage_count + (uint32_t)p, p);
        } else if (warp == 1) {
            // the whole warp walks the loop (every lane polls the barriers), one elected lane issues
            const bool lead = elect_one();
            if (tl && lead) tl[0] = clock64();
            mbar_wait(smem_u32(&ctrl->slab_full), pass_count & 1);
            tc_fence_after();
            if (tl && lead) tl[1] = clock64();
            long long waited = 0;
            const uint32_t idesc0 = make_idesc((uint32_t)items[0].n_cols);
            const uint32_t idesc1 = make_idesc((uint32_t)items[1].n_cols);
            uint32_t g = stage_count;
            for (int p = 0; p < A.n_pairs; ++p, ++g) {
                const uint32_t st = g % NSTAGE, ph = (g / NSTAGE) & 1;
                const long long w0 = tl ? clock64() : 0;
                mbar_wait(smem_u32(&ctrl->full[st]), ph);
                tc_fence_after();
                if (tl) waited += clock64() - w0;
                const uint32_t abase = b_addr + st * BSTAGE_BYTES;               // Toeplitz pair = A
#if MELF_TC_TMEM_A
                // the pair moves to tensor memory once (tcgen05.cp and tcgen05.mma run in issue
                // order, so neither the copy -> MMA nor the MMA -> next copy hazard needs a barrier)
                // and both items read it from there: shared memory then only streams the slabs
                const uint32_t ta = tmem + TMEM_A_COL + (g & 1u) * TMEM_A_COLS;
                if (lead) {
#pragma unroll
                    for (int ks = 0; ks < KS / 2; ++ks)
                        tmem_cp_128x256b(ta + 8 * ks, umma_desc(abase + ks * 2 * BSTRIP_BYTES, BSTRIP_BYTES));
                    umma_commit(smem_u32(&ctrl->empty[st]));
                }
#endif
                if (lead) {
#pragma unroll
                for (int slot = 0; slot < 2; ++slot) {
                    if (!items[slot].valid) continue;
                    const uint32_t bbase = slab_addr + slot * SLAB_BYTES + (2 * p) * 16;   // slab rows 2p.. = B
                    const uint32_t d = tmem + (uint32_t)(slot * NCOLS);
#pragma unroll
                    for (int ks = 0; ks < KS / 2; ++ks)
#if MELF_TC_TMEM_A
                        umma_i8_ta(d, ta + 8 * ks,
                                   umma_desc(bbase + ks * 2 * STRIP_BYTES, STRIP_BYTES),
                                   slot ? idesc1 : idesc0, (p > 0 || ks > 0) ? 1u : 0u);
#else
                        umma_i8(d, umma_desc(abase + ks * 2 * BSTRIP_BYTES, BSTRIP_BYTES),
                                umma_desc(bbase + ks * 2 * STRIP_BYTES, STRIP_BYTES),
                                slot ? idesc1 : idesc0, (p > 0 || ks > 0) ? 1u : 0u);
#endif
                }
#if !MELF_TC_TMEM_A
                umma_commit(smem_u32(&ctrl->empty[st]));
#endif
                }
                __syncwarp();
            }
            if (lead) umma_commit(smem_u32(&ctrl->accum));
            if (tl && lead) { tl[2] = clock64(); tl[7] = waited; }
        }
        stage_count += (uint32_t)A.n_pairs;
        __syncwarp();

        // ---- epilogue set-up (runs while the MMAs do) -----------------------------------------
        // warp -> TMEM lane quadrant qd (hardware rule: warp % 4), item slot, half of the columns.
        // Lane pair (2j, 2j+1) holds D[2x][.] and D[2x+1][.] of result column x = 16 qd + j; of
        // every 16-row chunk the even lane scores rows 0..7 and the odd lane rows 8..15, so each
        // needs 8 window sums per chunk: all of a warp's window sums are fetched up front.
        const int qd = warp & 3, slot = (warp >> 2) & 1, half = warp >> 3;
        const Item it = slot ? items[1] : items[0];
        const int n_chunks = it.n_cols >> 4;
        const int c_mid = (n_chunks + 1) >> 1;
        const int c0 = half ? c_mid : 0, c1 = half ? n_chunks : c_mid;      // <= 5 chunks
        const int x = 16 * qd + (lane >> 1);                       // result column inside the item
        const int odd = lane & 1;
        const bool col_ok = it.valid && x < it.w_c;
        const uint32_t *S = A.wsum + (size_t)(2 * pass + slot) * WSUM_WORDS + (size_t)x * WS_PITCH + 8 * odd;
        uint4 sw[5][2];
#pragma unroll
        for (int k = 0; k < 5; ++k)
            if (c0 + k < c1 && col_ok) {
                sw[k][0] = __ldg(reinterpret_cast<const uint4 *>(S + 16 * (c0 + k)));
                sw[k][1] = __ldg(reinterpret_cast<const uint4 *>(S + 16 * (c0 + k)) + 1);
            }

        mbar_wait(smem_u32(&ctrl->accum), pass_count & 1);
        tc_fence_after();
        if (tl && tid == 64) tl[3] = clock64();
        // every MMA of this pass has completed: slabs and Toeplitz ring are free again
        if (tid == 0 && has_next) {
            load_slabs(pass + gridDim.x);
            for (int p = 0; p < n_pre; ++p) load_stage(stage_count + (uint32_t)p, p);
        }
        __syncwarp();

        // ---- drain: numerators + arg-max -------------------------------------------------------
        const uint32_t taddr = tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(slot * NCOLS);
        long long best_num = 0;
        uint32_t best_idx = 0;
        bool has = false;
        float *map = WRITE_MAP ? A.result + (size_t)it.frame * A.geom.res_w * A.geom.res_h : nullptr;
        if (it.valid) {
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const int c = c0 + k;
                if (c < c1) {                                // warp-uniform
                    // 17 columns y0 .. y0+16 of this lane's row: even lanes use D[2x][y0 + i] = d[i],
                    // odd lanes D[2x+1][y0 + 1 + i] = d[i + 1]
                    uint32_t d16[16], d[17];
                    tmem_ld16(taddr + 16 * c, d16);
                    tmem_ld1(taddr + 16 * c + 16, d[16]);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) d[i] = d16[i];
                    const uint32_t swv[8] = {sw[k][0].x, sw[k][0].y, sw[k][0].z, sw[k][0].w,
                                             sw[k][1].x, sw[k][1].y, sw[k][1].z, sw[k][1].w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        // even sends D[2x][y0+8+j] and receives D[2x+1][y0+j+1]; odd the other way round
                        const uint32_t other = __shfl_xor_sync(0xffffffffu, odd ? d[j + 1] : d[8 + j], 1);
                        const uint32_t mine = odd ? d[9 + j] : d[j];
                        const int yl = 16 * c + 8 * odd + j;
                        if (col_ok && yl < it.h_c) {
                            const long long sum_tl = (long long)(int)(mine + other);
                            const long long num = A.n_t * sum_tl - A.sum_t * (long long)swv[j];
                            const uint32_t idx = (uint32_t)(it.y0 + yl) * (uint32_t)A.geom.res_w + (uint32_t)(it.x0 + x);
                            if (WRITE_MAP) map[idx] = score_of(num, A.n_t);
                            // this lane visits raster indices in increasing order, so a candidate replaces
                            // the running best only if its float32 score is strictly greater; scores are a
                            // monotone function of the numerator, and more than 4 float32 ulps of numerator
                            // distance ((|num| + n_t) >> 21) always separates the rounded values
                            if (!has) { best_num = num; best_idx = idx; has = true; }
                            else if (num > best_num) {
                                const long long mag = (best_num < 0 ? -best_num : best_num) + A.n_t;
                                if (num - best_num > (mag >> 21) + 1 || score_of(num, A.n_t) > score_of(best_num, A.n_t)) {
                                    best_num = num; best_idx = idx;
                                }
                            }
                        }
                    }
                }
            }
        }
        if (tl && tid == 64) tl[4] = clock64();
        u64 best = has ? melf_score_key(score_of(best_num, A.n_t), best_idx) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if (lane == 0) ctrl->red[pass_count & 1][warp] = best;
        if (tl && tid == 64) tl[5] = clock64();
        tc_fence_before();
        __syncthreads();            // TMEM fully read before the next pass's first MMA overwrites it
        tc_fence_after();
        if (tl && tid == 64) tl[6] = clock64();
        if (tid < 2) {
            // warps of slot s: 4 s .. 4 s + 3 and 8 + 4 s .. 8 + 4 s + 3
            const Item mine = tid ? items[1] : items[0];
            u64 b = 0;
            for (int w8 = 0; w8 < 8; ++w8) {
                const u64 k = ctrl->red[pass_count & 1][(w8 & 3) + 4 * tid + 8 * (w8 >> 2)];
                b = k > b ? k : b;
            }
            if (mine.valid && b) atomicMax(A.match_key + mine.frame, b);
        }
        __syncwarp();               // red[] alternates per pass: no thread can be two passes behind
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

// =======================================================================================
// Fused variant (default): ONE persistent kernel.  The operand preparation of the two-kernel
// version above runs in the shadow of the MMAs instead of in front of them: while warp 1
// streams the 960 tcgen05.mma of a pass (~82 k clk during which the other 14 warps used to
// idle), warps 2..15
//   (b) compute the window sums of the CURRENT pass's items from the lightness slabs that
//       already sit in shared memory -> a per-CTA global scratch (L2), read back in the drain;
//   (a) convert the BGR crop of the NEXT pass's items to lightness slabs -> per-CTA global
//       scratch (L2), from where the producer fetches them with one bulk copy each as soon as
//       the accumulators are complete.
// Scratch is 204 KB per CTA (30 MB per launch, L2-resident) instead of 99 KB per frame, so the
// only DRAM traffic left is the one read of the meter_rect crop.
// =======================================================================================
namespace {

#ifndef MELF_TC_EXPERIMENT
#define MELF_TC_EXPERIMENT 0    // measurement only (profiles/r02_timeline_z_*.txt): 1 extra commit, 2 extra fence, 3 extra barrier poll per stage
#endif
#ifndef MELF_DRAIN_WIDE
#define MELF_DRAIN_WIDE 0     // 1: 32 accumulator columns per tcgen05.ld, no prefetch -- measured no faster (drain 19-23 k clk per
#endif                        // pass against 17-21 k, profiles/r02_timeline_b_*.txt): the drain is not bound by the number of TMEM loads
#ifndef MELF_STAGE_RING
#define MELF_STAGE_RING 0     // 1: BGR rows reach the workers through a cp.async ring in shared memory -- measured SLOWER (staging
#endif                        // 55 k clk per pass against 48 k, profiles/r02_timeline_c_*.txt): global-load latency is not what binds it
#ifndef MELF_STAGE_TMA
#define MELF_STAGE_TMA 1         // frame bytes brought in by bulk copies (TMA), one 208-byte row segment each, per-warp buffers
#endif
#ifndef MELF_STAGE_TMA_FENCE
#define MELF_STAGE_TMA_FENCE 0   // 1: fence.proxy.async before every refill of a staging buffer (0: the loads out of it are forced to return first)
#endif
#ifndef MELF_STAGE_COALESCED
#define MELF_STAGE_COALESCED 0   // 1: the same buffers filled by cp.async, lane j = chunk j -- measured no faster than stage_slabs (LDGSTS: one request per lane per cycle)
#endif
#ifndef MELF_STAGE_L2_PREFETCH
#define MELF_STAGE_L2_PREFETCH 0   // 1: crop rows of the next pass prefetched into L2 before the window sums of this one -- measured
#endif                             // no gain (staging 46 k clk instead of 49 k, but the pass 96 k instead of 91 k: profiles/r02_timeline_e_*.txt)
#ifndef MELF_F_NSTAGE
#define MELF_F_NSTAGE 2
#define MELF_F_KSTEPS 8
#endif
constexpr int F_NSTAGE = MELF_F_NSTAGE;                       // Toeplitz ring
constexpr int F_KSTEPS = MELF_F_KSTEPS;                       // K-steps (of 32) per stage
constexpr int F_STAGE_BYTES = 2 * F_KSTEPS * BSTRIP_BYTES;    // 4096 per K-step
constexpr int F_SMEM_B = 2 * SLAB_BYTES;                      // 135168
constexpr int F_SMEM_H = F_SMEM_B + F_NSTAGE * F_STAGE_BYTES; // H table, 256 rows x 16 words, + segment totals
// behind the H table: the rest of the 227 KB, together the staging ring of the BGR rows (the window
// sums and the staging are consecutive phases of the same warps, so they can share the bytes)
constexpr int F_RING_AREA = 232448 - 512 - F_SMEM_H;          // 31232
constexpr int F_SMEM_CTRL = F_SMEM_H + F_RING_AREA;
constexpr int F_SMEM_TOTAL = F_SMEM_CTRL + 512;
constexpr int RING_ROWS = 16;                                 // slab rows per ring buffer
constexpr int RING_PITCH = 784;                               // 750-byte rows at any 16-byte phase: 49 chunks
constexpr int RING_CHUNKS = RING_PITCH / 16;
constexpr int RING_BUF_BYTES = RING_ROWS * RING_PITCH;        // 12544
static_assert(2 * RING_BUF_BYTES <= F_RING_AREA && 16384 + 4096 <= F_RING_AREA, "ring / H table do not fit");
constexpr int F_WORKERS = TC_THREADS - 64;                    // warps 2..15
constexpr int F_WORKER_WARPS = F_WORKERS / 32;
constexpr size_t F_SCRATCH_PER_CTA = 2 * (size_t)SLAB_BYTES + 2 * (size_t)WSUM_WORDS * 4;
constexpr size_t F_SCRATCH_PER_CTA_OV = 2 * (size_t)SLAB_BYTES + 4 * (size_t)WSUM_WORDS * 4;   // window sums per pass parity

struct CtrlF {
    unsigned long long full[F_NSTAGE];
    unsigned long long empty[F_NSTAGE];
    unsigned long long accum;
    unsigned long long slab_full;
    unsigned long long prep_done;
    uint32_t tmem_base;
    uint32_t pad;
    u64 red[2][16];
};

__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void worker_sync()
{
    asm volatile("bar.sync 1, %0;" :: "n"(F_WORKERS) : "memory");
}
__device__ __forceinline__ void fence_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// L2 residency: the frames stream through L2 once (evict-first), the per-CTA scratch is written
// and read back every pass and must stay there (evict-last); without the hints ncu shows the
// scratch bouncing through DRAM (profiles/r01_v7_ncu_match_fused.txt: 1.25 GB per launch for
// 0.77 GB of crops)
__device__ __forceinline__ u64 policy_evict_first()
{
    u64 p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ u64 policy_evict_last()
{
    u64 p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
#ifndef MELF_STAGE_PREFETCH
#define MELF_STAGE_PREFETCH 1
#endif
#ifndef MELF_HINT_STREAM
#define MELF_HINT_STREAM 2      // 1 (L1::no_allocate + L2 evict-first) measured 0.85 -> 1.06 ms: the misaligned 64-byte
#endif                          // spans of neighbouring lanes overlap and want L1; 2 = evict-first with L1 allocation
#ifndef MELF_HINT_KEEP
#define MELF_HINT_KEEP 1
#endif
__device__ __forceinline__ uint4 ldg_stream(const uint4 *ptr, u64 pol)
{
#if !MELF_HINT_STREAM
    return __ldg(ptr);
#endif
    uint4 v;
#if MELF_HINT_STREAM == 2
    asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(ptr), "l"(pol));
#else
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(ptr), "l"(pol));
#endif
    return v;
}
__device__ __forceinline__ void stg_keep(uint4 *ptr, uint4 v, u64 pol)
{
#if !MELF_HINT_KEEP
    *ptr = v; return;
#endif
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;"
                 :: "l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}
__device__ __forceinline__ void stg_keep2(uint2 *ptr, uint2 v, u64 pol)
{
#if !MELF_HINT_KEEP
    *ptr = v; return;
#endif
    asm volatile("st.global.L2::cache_hint.v2.u32 [%0], {%1, %2}, %3;"
                 :: "l"(ptr), "r"(v.x), "r"(v.y), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_load_keep(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, u64 pol)
{
#if !MELF_HINT_KEEP
    bulk_load(dst, src, bytes, bar); return;
#endif
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
}

// Lightness of 16 consecutive pixels (48 bytes = words v[0..11]) -> 16 bytes.
// max / min of (B, G, R) are taken on packed 16-bit lanes (VIMNMX3.U16x2: one instruction per
// pixel pair) with the colour byte in the HIGH byte of each lane -- the low byte is whatever the
// byte permute left there and can only break ties between equal colour bytes.  The float part
// is melf_lightness()'s operation order, two pixels per packed f32x2 instruction:
//     vmax = fl(max * k), vmin = fl(min * k)   as fma(2^23 + v, k, -2^23 k): exact product, one rounding
//     L = rint(fl(fl(vmax + vmin) * 127.5))    (x * .5) * 255 == x * 127.5; rint = + 1.5 * 2^23
// P = byte at which the first pixel starts inside v[0] (0 for callers that have aligned the words themselves)
template <int P, int NW>
__device__ __forceinline__ uint4 lightness16_at(const uint32_t (&v)[NW])
{
    const float2 k = melf_f2(1.f / 255.f), c = melf_f2(-8388608.f * (1.f / 255.f));
    const float2 half255 = melf_f2(127.5f), M = melf_f2(MELF_MAGIC_RINT);
    uint32_t o[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        uint32_t lb[4];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int b0 = P + 12 * g + 6 * q;               // byte index of B of the first pixel of the pair
            uint32_t ch[3];
#pragma unroll
            for (int cidx = 0; cidx < 3; ++cidx) {
                // bytes (b0 + cidx) and (b0 + cidx + 3) -> bytes 1 and 3
                const int ba = b0 + cidx, bb = ba + 3;
                const int wa = ba >> 2, wb = bb >> 2;        // wb == wa or wa + 1
                const uint32_t sa = (uint32_t)(ba & 3), sb = (uint32_t)((bb & 3) + (wb == wa ? 0 : 4));
                ch[cidx] = __byte_perm(v[wa], v[wb == wa ? wa : wb], (sb << 12) | (sb << 8) | (sa << 4) | sa);
            }
            const uint32_t mx = __vmaxu2(__vmaxu2(ch[0], ch[1]), ch[2]);
            const uint32_t mn = __vminu2(__vminu2(ch[0], ch[1]), ch[2]);
            // byte 1 / byte 3 -> 2^23 + value as float32
            const float2 hi = make_float2(__uint_as_float(__byte_perm(mx, 0x4B000000u, 0x7541u)),
                                          __uint_as_float(__byte_perm(mx, 0x4B000000u, 0x7543u)));
            const float2 lo = make_float2(__uint_as_float(__byte_perm(mn, 0x4B000000u, 0x7541u)),
                                          __uint_as_float(__byte_perm(mn, 0x4B000000u, 0x7543u)));
            const float2 sm = melf_add2(melf_fma2(hi, k, c), melf_fma2(lo, k, c));
            const float2 Lf = melf_add2_after_mul(melf_mul2(sm, half255), M);
            lb[2 * q] = __float_as_uint(Lf.x);
            lb[2 * q + 1] = __float_as_uint(Lf.y);
        }
        const uint32_t lo2 = __byte_perm(lb[0], lb[1], 0x0040), hi2 = __byte_perm(lb[2], lb[3], 0x0040);
        o[g] = __byte_perm(lo2, hi2, 0x5410);
    }
    return make_uint4(o[0], o[1], o[2], o[3]);
}
__device__ __forceinline__ uint4 lightness16(const uint32_t (&v)[12]) { return lightness16_at<0, 12>(v); }
// the same for 16 words holding the 48 bytes from byte `p` (0..15) on: the shift is folded into the byte
// selectors, one instance per phase behind a warp-uniform switch
__device__ __forceinline__ uint4 lightness16_phase(const uint32_t (&e)[16], uint32_t p)
{
    switch (p) {
    case 0: return lightness16_at<0, 16>(e);
    case 1: return lightness16_at<1, 16>(e);
    case 2: return lightness16_at<2, 16>(e);
    case 3: return lightness16_at<3, 16>(e);
    case 4: return lightness16_at<4, 16>(e);
    case 5: return lightness16_at<5, 16>(e);
    case 6: return lightness16_at<6, 16>(e);
    case 7: return lightness16_at<7, 16>(e);
    case 8: return lightness16_at<8, 16>(e);
    case 9: return lightness16_at<9, 16>(e);
    case 10: return lightness16_at<10, 16>(e);
    case 11: return lightness16_at<11, 16>(e);
    case 12: return lightness16_at<12, 16>(e);
    case 13: return lightness16_at<13, 16>(e);
    case 14: return lightness16_at<14, 16>(e);
    default: return lightness16_at<15, 16>(e);
    }
}

// 16 consecutive pixels of slab row r / strip s of an item (fused kernel's version of stage_unit)
__device__ __forceinline__ uint4 stage_unit_f(const MelfMatchTcArgs &A, const Item &it, int r, int s, u64 pol_stream)
{
    const int cy = it.y0 + r, cx = it.x0 + 16 * s;
    if (cy >= A.geom.crop_h || cx >= A.geom.crop_w) return make_uint4(0, 0, 0, 0);
    const uint8_t *src = A.img + (size_t)it.frame * A.frame_stride +
                         (size_t)(A.geom.crop_y0 + cy) * A.row_stride +
                         (size_t)A.channels * (size_t)(A.geom.crop_x0 + cx);
    const uintptr_t addr = reinterpret_cast<uintptr_t>(src);
    const uint4 *al = reinterpret_cast<const uint4 *>(addr & ~(uintptr_t)15);
    const uint32_t off = (uint32_t)(addr & 15);
    const uint32_t ws = off >> 2, bs = (off & 3) * 8;
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
    const int nchunk = A.channels == 3 ? 4 : 2;
    uint32_t e[16];
#pragma unroll
    for (int cidx = 0; cidx < 4; ++cidx) {
        uint4 q = make_uint4(0, 0, 0, 0);
        if (A.debug_skip & 16) {
            // measurement only: the conversion without its global loads
            q = make_uint4((uint32_t)addr * 2654435761u + cidx, (uint32_t)addr ^ 0x9e3779b9u, (uint32_t)(addr >> 3) + 77u * cidx, (uint32_t)addr * 40503u);
        } else
        if (cidx < nchunk && reinterpret_cast<uintptr_t>(al + cidx) < limit) q = ldg_stream(al + cidx, pol_stream);
        e[4 * cidx] = q.x; e[4 * cidx + 1] = q.y; e[4 * cidx + 2] = q.z; e[4 * cidx + 3] = q.w;
    }
    if (A.debug_skip & 32)      // measurement only: the global loads without the conversion
        return make_uint4(e[0] ^ e[4] ^ e[8] ^ e[12], e[1] ^ e[5] ^ e[9] ^ e[13], e[2] ^ e[6] ^ e[10] ^ e[14], e[3] ^ e[7] ^ e[11] ^ e[15]);
    uint32_t v[12];
    // shift right by `off` bytes; the word part is uniform across a warp whenever the row stride
    // is a multiple of 16, so the branches normally do not diverge
    switch (ws) {
    case 0:
#pragma unroll
        for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j], e[j + 1], bs);
        break;
    case 1:
#pragma unroll
        for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 1], e[j + 2], bs);
        break;
    case 2:
#pragma unroll
        for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 2], e[j + 3], bs);
        break;
    default:
#pragma unroll
        for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 3], j + 4 < 16 ? e[j + 4] : 0u, bs);
        break;
    }
    if (A.channels == 1) return make_uint4(v[0], v[1], v[2], v[3]);
    return lightness16(v);
}

// Window sums of one item, S(y, x) = sum_{i<th} sum_{j<tw} L(y + i, x + j), from the slab in
// shared memory to global scratch (y-major [y][64]); called by all F_WORKERS threads.
// Two x-halves of 32 columns, each in three steps separated by worker barriers:
//   H   thread = slab row: horizontal sums H(row, x) sliding along x (the start value of the
//       second half is carried in a register), u16 pairs, word (row, xp) at row*16 + (xp ^ (row & 15))
//   V1  thread = (8-row segment, x pair): segment totals C[seg][xp]
//   V2  thread = (8 result rows, x pair): first window = whole segments + (th & 7) single rows,
//       then sliding down its 8 rows
__device__ __forceinline__ void window_sums(const MelfMatchTcArgs &A, const Item &it, const uint8_t *slab,
                                            uint32_t *Htab, uint32_t *S, int wt)
{
    uint2 *Ctab = reinterpret_cast<uint2 *>(Htab + 256 * 16);          // [32 segments][16 x pairs]
    const u64 pol_keep = policy_evict_last();
    const int n_rows = min(SLAB_DATA_ROWS, it.h_c + A.tmpl_h);          // H rows that are read
    const int row = wt;                                                 // n_rows <= 256 < F_WORKERS
    const uint8_t *sl = slab + row * 16;
    const uint32_t o = (uint32_t)A.tmpl_w & 15, ws = o >> 2, bs = (o & 3) * 8;
    int h = 0;
    if (row < n_rows) {
#pragma unroll
        for (int s = 0; s < KS; ++s) {
            const uint4 q = *reinterpret_cast<const uint4 *>(sl + s * STRIP_BYTES);
            h = (int)__dp4a(q.x, A.ones_words[4 * s], (uint32_t)h);
            h = (int)__dp4a(q.y, A.ones_words[4 * s + 1], (uint32_t)h);
            h = (int)__dp4a(q.z, A.ones_words[4 * s + 2], (uint32_t)h);
            h = (int)__dp4a(q.w, A.ones_words[4 * s + 3], (uint32_t)h);
        }
    }
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
        if (row < n_rows) {
            // leaving bytes: columns 32 half .. +31; entering bytes: columns tw + 32 half .. +31
            uint32_t w[8], e[12], g[8];
#pragma unroll
            for (int s = 0; s < 2; ++s) {
                const uint4 q = *reinterpret_cast<const uint4 *>(sl + (2 * half + s) * STRIP_BYTES);
                w[4 * s] = q.x; w[4 * s + 1] = q.y; w[4 * s + 2] = q.z; w[4 * s + 3] = q.w;
            }
            const int s0 = (A.tmpl_w >> 4) + 2 * half;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                uint4 q = make_uint4(0, 0, 0, 0);
                if (s0 + s < KS) q = *reinterpret_cast<const uint4 *>(sl + (s0 + s) * STRIP_BYTES);
                e[4 * s] = q.x; e[4 * s + 1] = q.y; e[4 * s + 2] = q.z; e[4 * s + 3] = q.w;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t t0 = ws == 0 ? e[j] : ws == 1 ? e[j + 1] : ws == 2 ? e[j + 2] : e[j + 3];
                const uint32_t t1 = ws == 0 ? e[j + 1] : ws == 1 ? e[j + 2] : ws == 2 ? e[j + 3] : e[j + 4];
                g[j] = __funnelshift_r(t0, t1, bs);
            }
            uint32_t *Hb = Htab + row * 16;
            const int sw = row & 15;
#pragma unroll
            for (int xp = 0; xp < 16; ++xp) {
                const int xa = 2 * xp, xb = 2 * xp + 1;
                const uint32_t lo = (uint32_t)h;
                h += (int)((g[xa >> 2] >> (8 * (xa & 3))) & 255) - (int)((w[xa >> 2] >> (8 * (xa & 3))) & 255);
                const uint32_t hi = (uint32_t)h;
                h += (int)((g[xb >> 2] >> (8 * (xb & 3))) & 255) - (int)((w[xb >> 2] >> (8 * (xb & 3))) & 255);
                Hb[xp ^ sw] = lo | (hi << 16);
            }
        }
        worker_sync();
        // V1: totals of the 8-row segments (rows beyond n_rows are never part of a window that is used)
        for (int t = wt; t < 32 * 16; t += F_WORKERS) {
            const int seg = t >> 4, xp = t & 15;
            uint32_t c0 = 0, c1 = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = 8 * seg + i;
                const uint32_t q = Htab[r * 16 + (xp ^ (r & 15))];
                c0 += q & 0xffff; c1 += q >> 16;
            }
            Ctab[t] = make_uint2(c0, c1);
        }
        worker_sync();
        // V2
        {
            const int yseg = wt >> 4, xp = wt & 15;
            const int ya = 8 * yseg, yb = min(it.h_c, ya + 8);
            if (ya < yb) {
                uint32_t s0 = 0, s1 = 0;
                const int n_seg = A.tmpl_h >> 3;
                for (int k = 0; k < n_seg; ++k) {
                    const uint2 c = Ctab[(yseg + k) * 16 + xp];
                    s0 += c.x; s1 += c.y;
                }
                for (int r = ya + 8 * n_seg; r < ya + A.tmpl_h; ++r) {
                    const uint32_t q = Htab[r * 16 + (xp ^ (r & 15))];
                    s0 += q & 0xffff; s1 += q >> 16;
                }
                for (int y = ya; y < yb; ++y) {
                    stg_keep2(reinterpret_cast<uint2 *>(S + y * 64 + 32 * half + 2 * xp), make_uint2(s0, s1), pol_keep);
                    const int rin = y + A.tmpl_h;                                  // <= 255 (plan cap)
                    const uint32_t vin = Htab[rin * 16 + (xp ^ (rin & 15))];
                    const uint32_t vout = Htab[y * 16 + (xp ^ (y & 15))];
                    s0 += (vin & 0xffff) - (vout & 0xffff);
                    s1 += (vin >> 16) - (vout >> 16);
                }
            }
        }
        worker_sync();                          // H table free again
    }
}

// The crop rows of the items of a LATER pass pulled from DRAM into L2 (one bulk prefetch per row), issued a
// whole window-sum phase before stage_slabs reads them: ncu's source view had 25 % of all stall samples of
// the kernel on the first use of stage_unit_f's global loads (long scoreboard, profiles/r02_ncu_locate_straight_source.txt)
// -- DRAM latency with 14 warps and one unit in flight per lane.  Called by all F_WORKERS threads.
__device__ __forceinline__ void prefetch_items_l2(const MelfMatchTcArgs &A, const Item (&its)[2], int wt)
{
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
        const Item it = slot ? its[1] : its[0];
        if (!it.valid) continue;
        const int n_rows = min(SLAB_DATA_ROWS, A.geom.crop_h - it.y0);
        const int row_bytes = A.channels * min(KS * 16, A.geom.crop_w - it.x0);
        const uint8_t *base = A.img + (size_t)it.frame * A.frame_stride + (size_t)(A.geom.crop_y0 + it.y0) * A.row_stride +
                              (size_t)A.channels * (size_t)(A.geom.crop_x0 + it.x0);
        for (int r = wt; r < n_rows; r += F_WORKERS) {
            const uintptr_t src = reinterpret_cast<uintptr_t>(base + (size_t)r * A.row_stride);
            const uintptr_t a16 = src & ~(uintptr_t)15;
            uintptr_t end = (src + (uintptr_t)row_bytes + 15) & ~(uintptr_t)15;
            if (end > limit) end = limit;
            if (end > a16)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(a16), "r"((uint32_t)(end - a16)) : "memory");
        }
    }
}

// lightness slabs of the (up to two) items of one pass -> this CTA's global scratch
__device__ __forceinline__ void stage_slabs(const MelfMatchTcArgs &A, const Item (&its)[2], uint8_t *scratch,
                                            int wwarp, int lane, int n_wwarps = F_WORKER_WARPS)
{
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
        const Item it = slot ? its[1] : its[0];
        if (!it.valid) continue;
        const int n_rows = min(SLAB_DATA_ROWS, A.geom.crop_h - it.y0);
        const int n_units = ((n_rows + 7) >> 3) * 4;
        uint8_t *dst = scratch + (size_t)slot * SLAB_BYTES;
        const u64 pol_stream = policy_evict_first(), pol_keep = policy_evict_last();
        // warp-unit = 8 rows x 4 strips; lane = strip * 8 + row: every quarter warp stores 128
        // contiguous bytes and reads 8 rows x 192 contiguous bytes of the frame
        for (int u = wwarp; u < n_units; u += n_wwarps) {
            const int rg = u >> 2, sg = u & 3;
            const int r = rg * 8 + (lane & 7), s = sg * 4 + (lane >> 3);
#if MELF_STAGE_PREFETCH
            {
                // pull the next unit's bytes into L1 while this one is converted (no registers held)
                const int un = u + n_wwarps;
                const int rn = (un >> 2) * 8 + (lane & 7), sn = (un & 3) * 4 + (lane >> 3);
                const int cy = it.y0 + rn, cx = it.x0 + 16 * sn;
                if (un < n_units && cy < A.geom.crop_h && cx < A.geom.crop_w) {
                    const uint8_t *src = A.img + (size_t)it.frame * A.frame_stride +
                                         (size_t)(A.geom.crop_y0 + cy) * A.row_stride +
                                         (size_t)A.channels * (size_t)(A.geom.crop_x0 + cx);
                    asm volatile("prefetch.global.L1 [%0];" :: "l"(src));
                    asm volatile("prefetch.global.L1 [%0];" :: "l"(src + 16 * A.channels - 1));
                }
            }
#endif
            stg_keep(reinterpret_cast<uint4 *>(dst + s * STRIP_BYTES + r * 16), stage_unit_f(A, it, r, s, pol_stream), pol_keep);
        }
    }
}

// stage_slabs with COALESCED frame loads.  In stage_slabs a lane fetches its own 64-byte window (row = lane & 7,
// strip = lane >> 3): one LDG.128 of a warp touches 8 rows x 2-3 cache lines and the four of them touch the same
// lines again -- about 80 tag look-ups per 1.5 KB unit.  The skip experiments (profiles/r02_skip_experiments_g.txt)
// show the phase bound by exactly these loads (47 k clk per pass with the conversion removed, 28 k with the loads
// removed) and slowed further by the MMAs' operand traffic on the same L1 / shared-memory data path.  Here the
// warp reads the 8 row segments of a unit as consecutive 16-byte chunks (lane j = chunk j: about 7 lines per
// LDG.128), parks them in a private 1664-byte shared-memory buffer and every lane picks its window from there
// (row pitch 208 B = 13 chunks: the eight rows of a quarter warp fall into eight different bank groups).
constexpr int CO_ROW_CHUNKS = 13;                             // 192 bytes at any 16-byte phase
constexpr int CO_BUF_BYTES = 8 * CO_ROW_CHUNKS * 16;          // 1664 per worker warp
static_assert(F_WORKER_WARPS * CO_BUF_BYTES <= F_RING_AREA, "per-warp staging buffers do not fit");

__device__ __forceinline__ void stage_slabs_co(const MelfMatchTcArgs &A, const Item (&its)[2], uint8_t *scratch,
                                               uint8_t *wbuf, int wwarp, int lane)
{
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
    const int row_chunks = A.channels == 3 ? CO_ROW_CHUNKS : 5;            // 64 pixels of a row segment + phase
    const int n_chunks = 8 * row_chunks;
    const int seg_bytes = 64 * A.channels;
    const uint32_t wbuf_s = smem_u32(wbuf);
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
        const Item it = slot ? its[1] : its[0];
        if (!it.valid) continue;
        const int n_rows = min(SLAB_DATA_ROWS, A.geom.crop_h - it.y0);
        const int n_units = ((n_rows + 7) >> 3) * 4;
        uint8_t *dst = scratch + (size_t)slot * SLAB_BYTES;
        const u64 pol_keep = policy_evict_last();
        const uint8_t *base = A.img + (size_t)it.frame * A.frame_stride + (size_t)(A.geom.crop_y0 + it.y0) * A.row_stride +
                              (size_t)A.channels * (size_t)(A.geom.crop_x0 + it.x0);
        // the unit's bytes, chunk by chunk, global -> shared without passing registers (cp.async: all four chunks
        // of a lane in flight together; with LDG + STS ptxas, at the 128-register cap, reused the destination
        // registers and serialised the round trips)
        auto issue = [&](int u) {
            const int rg = u >> 2, sg = u & 3;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = 32 * c + lane;
                if (j < n_chunks) {
                    const int rr = A.channels == 3 ? j / CO_ROW_CHUNKS : j / 5;
                    const int cc = j - rr * row_chunks;
                    const uintptr_t seg = reinterpret_cast<uintptr_t>(base + (size_t)(rg * 8 + rr) * A.row_stride) +
                                          (uintptr_t)(seg_bytes * sg);
                    const uintptr_t a16 = (seg & ~(uintptr_t)15) + 16u * (uintptr_t)cc;
                    // rows behind the crop are never looked at; a chunk whose tail leaves the frame memory is cut (zero fill)
                    if (rg * 8 + rr < n_rows && a16 < limit) {
                        const uint32_t n_src = a16 + 16 <= limit ? 16u : (uint32_t)(limit - a16);
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;"
                                     :: "r"(wbuf_s + 16u * (uint32_t)j), "l"(a16), "r"(n_src) : "memory");
                    }
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        if (wwarp < n_units) issue(wwarp);
#pragma unroll 1
        for (int u = wwarp; u < n_units; u += F_WORKER_WARPS) {
            const int rg = u >> 2, sg = u & 3;
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            // ---- lane = (row, strip): its 64-byte window out of the buffer --------------------------------
            const int r = lane & 7, sl = lane >> 3;
            const int row = rg * 8 + r, sidx = sg * 4 + sl;
            const bool live = it.y0 + row < A.geom.crop_h && it.x0 + 16 * sidx < A.geom.crop_w;
            const uintptr_t seg = reinterpret_cast<uintptr_t>(base + (size_t)row * A.row_stride) + (uintptr_t)(seg_bytes * sg);
            const uint32_t pos = (uint32_t)(seg & 15) + (uint32_t)(16 * A.channels * sl);
            const uint8_t *rp = wbuf + r * (16 * row_chunks) + (pos & ~15u);
            uint32_t e[17];
#pragma unroll
            for (int cidx = 0; cidx < 4; ++cidx) {
                uint4 q = make_uint4(0, 0, 0, 0);
                if (live && (int)(pos >> 4) + cidx < row_chunks) q = *reinterpret_cast<const uint4 *>(rp + 16 * cidx);
                e[4 * cidx] = q.x; e[4 * cidx + 1] = q.y; e[4 * cidx + 2] = q.z; e[4 * cidx + 3] = q.w;
            }
            e[16] = 0;
            __syncwarp();
            // the buffer is free again: the next unit's bytes travel while this one is converted
            if (u + F_WORKER_WARPS < n_units) issue(u + F_WORKER_WARPS);
            uint4 out = make_uint4(0, 0, 0, 0);
            if (live) {
                const uint32_t ws = (pos & 15u) >> 2, bs = (pos & 3u) * 8;
                uint32_t v[12];
                switch (ws) {
                case 0:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j], e[j + 1], bs);
                    break;
                case 1:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 1], e[j + 2], bs);
                    break;
                case 2:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 2], e[j + 3], bs);
                    break;
                default:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 3], e[j + 4], bs);
                    break;
                }
                out = A.channels == 1 ? make_uint4(v[0], v[1], v[2], v[3]) : lightness16(v);
            }
            stg_keep(reinterpret_cast<uint4 *>(dst + sidx * STRIP_BYTES + row * 16), out, pol_keep);
        }
    }
}

// stage_slabs with the frame bytes brought in by the TMA engine: per unit (8 rows x 64 pixels) eight bulk
// copies of one 208-byte row segment each (lanes 0-7 issue one apiece) into the warp's private buffer, completion
// on the warp's own mbarrier; the next unit's copies are issued as soon as every lane has taken its 64-byte
// window out of the buffer, so they travel while this unit is converted.  Why: the load/store units of an SM
// serve a 16-byte-per-lane access at one 128-byte line per cycle and quarter warp, and stage_slabs' pattern
// (8 different rows per quarter warp) makes each of its four LDG.128 cost 32 of those cycles -- 128+ cycles per
// unit, the whole staging phase (profiles/r02_skip_experiments_g.txt: 47 k clk per pass with the conversion
// removed).  cp.async (LDGSTS) is no better, one request per lane per cycle (r02_timeline_h_*.txt).  Bulk copies
// do not pass the load/store units at all.
__device__ __forceinline__ void stage_slabs_tma(const MelfMatchTcArgs &A, const Item (&its)[2], uint8_t *scratch,
                                                uint8_t *wbuf, uint32_t bar, uint32_t &phase, int wwarp, int lane)
{
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
    const int row_chunks = A.channels == 3 ? CO_ROW_CHUNKS : 5;            // 64 pixels of a row segment + phase
    const int seg_bytes = 64 * A.channels;
    const uint32_t wbuf_s = smem_u32(wbuf);
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
        const Item it = slot ? its[1] : its[0];
        if (!it.valid) continue;
        const int n_rows = min(SLAB_DATA_ROWS, A.geom.crop_h - it.y0);
        const int n_units = ((n_rows + 7) >> 3) * 4;
        uint8_t *dst = scratch + (size_t)slot * SLAB_BYTES;
        const u64 pol_keep = policy_evict_last(), pol_stream = policy_evict_first();
        const uint8_t *base = A.img + (size_t)it.frame * A.frame_stride + (size_t)(A.geom.crop_y0 + it.y0) * A.row_stride +
                              (size_t)A.channels * (size_t)(A.geom.crop_x0 + it.x0);
        auto issue = [&](int u) {
            const int rg = u >> 2, sg = u & 3;
            uintptr_t a16 = 0;
            uint32_t bytes = 0;
            if (lane < 8 && rg * 8 + lane < n_rows) {
                const uintptr_t seg = reinterpret_cast<uintptr_t>(base + (size_t)(rg * 8 + lane) * A.row_stride) +
                                      (uintptr_t)(seg_bytes * sg);
                a16 = seg & ~(uintptr_t)15;
                if (a16 < limit) bytes = (uint32_t)min((uintptr_t)(16 * row_chunks), limit - a16);   // cut at the end of the frame memory
            }
            const uint32_t total = __reduce_add_sync(0xffffffffu, bytes);
#if MELF_STAGE_TMA_FENCE
            // the generic-proxy reads of the previous unit come before the async-proxy writes of this one
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
            if (lane == 0) mbar_expect_tx(bar, total);
            __syncwarp();
            // (frame bytes pass through L2 once: evict-first, so that they do not push out the scratch slabs)
            if (bytes) bulk_load_keep(wbuf_s + (uint32_t)(lane * 16 * row_chunks), reinterpret_cast<const void *>(a16), bytes, bar, pol_stream);
        };
        if (wwarp < n_units) issue(wwarp);
#pragma unroll 1
        for (int u = wwarp; u < n_units; u += F_WORKER_WARPS) {
            const int rg = u >> 2, sg = u & 3;
            mbar_wait(bar, phase & 1);
            ++phase;
            // ---- lane = (row, strip): its 64-byte window out of the buffer --------------------------------
            const int r = lane & 7, sl = lane >> 3;
            const int row = rg * 8 + r, sidx = sg * 4 + sl;
            const bool live = it.y0 + row < A.geom.crop_h && it.x0 + 16 * sidx < A.geom.crop_w;
            const uintptr_t seg = reinterpret_cast<uintptr_t>(base + (size_t)row * A.row_stride) + (uintptr_t)(seg_bytes * sg);
            const uint32_t pos = (uint32_t)(seg & 15) + (uint32_t)(16 * A.channels * sl);
            const uint8_t *rp = wbuf + r * (16 * row_chunks) + (pos & ~15u);
            uint32_t e[17];
#pragma unroll
            for (int cidx = 0; cidx < 4; ++cidx) {
                uint4 q = make_uint4(0, 0, 0, 0);
                if (live && (int)(pos >> 4) + cidx < row_chunks) q = *reinterpret_cast<const uint4 *>(rp + 16 * cidx);
                e[4 * cidx] = q.x; e[4 * cidx + 1] = q.y; e[4 * cidx + 2] = q.z; e[4 * cidx + 3] = q.w;
            }
            e[16] = 0;
#if !MELF_STAGE_TMA_FENCE
            {
                // the four loads must have RETURNED (not just been issued) before the buffer is handed back:
                // a real instruction that reads one register of each
                uint32_t seen;
                asm volatile("xor.b32 %0, %1, %2;\n\txor.b32 %0, %0, %3;\n\txor.b32 %0, %0, %4;"
                             : "=r"(seen) : "r"(e[0]), "r"(e[4]), "r"(e[8]), "r"(e[12]));
                (void)seen;
            }
#endif
            __syncwarp();
            // the buffer is free again: the next unit's bytes travel while this one is converted
            if (u + F_WORKER_WARPS < n_units) issue(u + F_WORKER_WARPS);
            uint4 out = make_uint4(0, 0, 0, 0);
            if (live) {
                const uint32_t ws = (pos & 15u) >> 2, bs = (pos & 3u) * 8;
                uint32_t v[12];
                switch (ws) {
                case 0:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j], e[j + 1], bs);
                    break;
                case 1:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 1], e[j + 2], bs);
                    break;
                case 2:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 2], e[j + 3], bs);
                    break;
                default:
#pragma unroll
                    for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 3], e[j + 4], bs);
                    break;
                }
                out = A.channels == 1 ? make_uint4(v[0], v[1], v[2], v[3]) : lightness16(v);
            }
            stg_keep(reinterpret_cast<uint4 *>(dst + sidx * STRIP_BYTES + row * 16), out, pol_keep);
        }
    }
}

// stage_slabs_tma with ONE tiled tensor copy per unit instead of eight row copies: the frame batch is described
// by a 3-D tensor map (bytes of a row, rows, frames; built per launch by melf_launch_match_fused_ov when the
// caller's base address and strides are multiples of 16), the box is 8 rows x (64 pixels + 16 bytes) from the
// 16-byte boundary at or before the unit's first pixel.  (The innermost start coordinate of a u8 tensor copy has
// to be a multiple of 16 BYTES: anything else ends the kernel with "illegal instruction" -- tools/tmap_probe.cu,
// measured on this B200; the 16 spare bytes of the box are for that.)  Against the row copies this removes, per
// unit, the eight-trip issue loop (a divergent UBLKCP is serialised lane by lane) and the twelve funnel shifts that
// realigned each lane's window: with 16-byte-multiple strides the byte phase is the same for every unit of the
// launch, so it is folded into the conversion's byte selectors (lightness16_phase).
// The 208-byte row pitch keeps the quarter-warp's eight rows on eight different 16-byte bank groups (13 r mod 8).
// Rows and bytes beyond the frame come back as zeros (tensor-map bounds), so there is no end-of-memory clamp.
__device__ __forceinline__ void tensor_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2,
                                               uint32_t bar, u64 pol)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
                 " [%0], [%1, {%2, %3, %4}], [%5], %6;"
                 :: "r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar), "l"(pol)
                 : "memory");
}

__device__ __forceinline__ void stage_slabs_tmap(const MelfMatchTcArgs &A, const CUtensorMap *map, const Item (&its)[2],
                                                 uint8_t *scratch, uint8_t *wbuf, uint32_t bar, uint32_t &phase,
                                                 int wwarp, int lane)
{
    const int pitch = 64 * A.channels + 16;                                // box row: 208 (BGR) / 80 bytes
    const uint32_t box_bytes = 8u * (uint32_t)pitch;
    const uint32_t wbuf_s = smem_u32(wbuf);
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
        const Item it = slot ? its[1] : its[0];
        if (!it.valid) continue;
        const int n_rows = min(SLAB_DATA_ROWS, A.geom.crop_h - it.y0);
        const int n_units = ((n_rows + 7) >> 3) * 4;
        uint8_t *dst = scratch + (size_t)slot * SLAB_BYTES;
        const u64 pol_keep = policy_evict_last(), pol_stream = policy_evict_first();
        // first byte of the item's rows; 64-pixel segments are whole 16-byte chunks, so `ph` holds for every unit
        const int cx0 = A.channels * (A.geom.crop_x0 + it.x0), cy0 = A.geom.crop_y0 + it.y0;
        const uint32_t ph = (uint32_t)cx0 & 15u;
        auto issue = [&](int u) {
            if (lane == 0) {
                mbar_expect_tx(bar, box_bytes);
                tensor_load_3d(wbuf_s, map, (cx0 & ~15) + 64 * A.channels * (u & 3), cy0 + 8 * (u >> 2), it.frame, bar, pol_stream);
            }
        };
        if (wwarp < n_units) issue(wwarp);
#pragma unroll 1
        for (int u = wwarp; u < n_units; u += F_WORKER_WARPS) {
            const int rg = u >> 2, sg = u & 3;
            mbar_wait(bar, phase & 1);
            ++phase;
            const int r = lane & 7, sl = lane >> 3;
            const int row = rg * 8 + r, sidx = sg * 4 + sl;
            const bool live = it.y0 + row < A.geom.crop_h && it.x0 + 16 * sidx < A.geom.crop_w;
            const uint8_t *rp = wbuf + r * pitch + 16 * A.channels * sl;
            const int n_chunks = A.channels == 3 ? 4 : 2;
            uint32_t e[16];
#pragma unroll
            for (int cidx = 0; cidx < 4; ++cidx) {
                uint4 q = make_uint4(0, 0, 0, 0);
                if (live && cidx < n_chunks) q = *reinterpret_cast<const uint4 *>(rp + 16 * cidx);
                e[4 * cidx] = q.x; e[4 * cidx + 1] = q.y; e[4 * cidx + 2] = q.z; e[4 * cidx + 3] = q.w;
            }
            {
                // the loads must have RETURNED (not just been issued) before the buffer is handed back
                uint32_t seen;
                asm volatile("xor.b32 %0, %1, %2;\n\txor.b32 %0, %0, %3;\n\txor.b32 %0, %0, %4;"
                             : "=r"(seen) : "r"(e[0]), "r"(e[4]), "r"(e[8]), "r"(e[12]));
                (void)seen;
            }
            __syncwarp();
            if (u + F_WORKER_WARPS < n_units) issue(u + F_WORKER_WARPS);
            uint4 out = make_uint4(0, 0, 0, 0);
            if (live) {
                if (A.channels == 3) {
                    out = lightness16_phase(e, ph);
                } else {
                    const uint32_t ws = ph >> 2, bs = (ph & 3u) * 8;
                    uint32_t v[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t lo = ws == 0 ? e[j] : ws == 1 ? e[j + 1] : ws == 2 ? e[j + 2] : e[j + 3];
                        const uint32_t hi = ws == 0 ? e[j + 1] : ws == 1 ? e[j + 2] : ws == 2 ? e[j + 3] : e[j + 4];
                        v[j] = __funnelshift_r(lo, hi, bs);
                    }
                    out = make_uint4(v[0], v[1], v[2], v[3]);
                }
            }
            stg_keep(reinterpret_cast<uint4 *>(dst + sidx * STRIP_BYTES + row * 16), out, pol_keep);
        }
    }
}

// stage_slabs with the frame bytes passing through shared memory: every group of RING_ROWS crop rows is
// fetched with cp.async (16-byte chunks, no registers held while the bytes are in flight) into one of two
// ring buffers while the previous group is converted out of the other.  The straight version waits ~2.6 k
// clk per unit on its global loads (24 % of all stall samples sat on their first use, profiles/
// r01_v9_ncu_match_fused.txt); here the conversion reads shared memory and the latency of group g + 1 hides
// behind the conversion of group g.  Called by all F_WORKERS threads.
__device__ __forceinline__ void stage_slabs_ring(const MelfMatchTcArgs &A, const Item (&its)[2], uint8_t *scratch,
                                                 uint8_t *ring, int wt)
{
    const uint32_t ring_s = smem_u32(ring);
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
        const Item it = slot ? its[1] : its[0];
        if (!it.valid) continue;
        const int n_rows = min(SLAB_DATA_ROWS, A.geom.crop_h - it.y0);
        const int n_groups = (n_rows + RING_ROWS - 1) / RING_ROWS;
        const int n_px = min(KS * 16, A.geom.crop_w - it.x0);                 // pixels of a row that a slab can hold
        const int row_bytes = A.channels * n_px;
        const uint8_t *base = A.img + (size_t)it.frame * A.frame_stride + (size_t)(A.geom.crop_y0 + it.y0) * A.row_stride +
                              (size_t)A.channels * (size_t)(A.geom.crop_x0 + it.x0);
        uint8_t *dst = scratch + (size_t)slot * SLAB_BYTES;
        const u64 pol_keep = policy_evict_last();
        auto issue = [&](int g) {
            const uint32_t buf = ring_s + (uint32_t)(g & 1) * RING_BUF_BYTES;
            for (int i = wt; i < RING_ROWS * RING_CHUNKS; i += F_WORKERS) {
                const int r = i / RING_CHUNKS, c = i - r * RING_CHUNKS;
                const int row = g * RING_ROWS + r;
                if (row >= n_rows) continue;
                const uintptr_t src = reinterpret_cast<uintptr_t>(base + (size_t)row * A.row_stride);
                const uintptr_t a16 = (src & ~(uintptr_t)15) + 16u * (uintptr_t)c;
                if (a16 >= src + (uintptr_t)row_bytes) continue;               // behind the last byte needed
                // bytes behind the end of the frame memory are not read (zero fill)
                // (the bytes needed lie inside the frame memory, the tail of their last chunk may not: zero fill)
                const uint32_t n_src = a16 + 16 <= limit ? 16u : (uint32_t)(limit - a16);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;"
                             :: "r"(buf + (uint32_t)(r * RING_PITCH + 16 * c)), "l"(a16), "r"(n_src) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        issue(0);
#pragma unroll 1
        for (int g = 0; g < n_groups; ++g) {
            if (g + 1 < n_groups) issue(g + 1); else asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            worker_sync();                                   // everybody's chunks of group g have landed
            if (wt < RING_ROWS * KS) {
                // unit = 16 pixels; lane = (strip & 3) * 8 + row & 7: a quarter warp stores 128 contiguous bytes
                const int r = (wt & 7) + 8 * (wt >> 7), sidx = (wt >> 3) & 15;
                const int row = g * RING_ROWS + r;
                if (row < ((n_rows + 7) & ~7)) {
                    uint4 out = make_uint4(0, 0, 0, 0);
                    if (row < n_rows && 16 * sidx < n_px) {
                        const uintptr_t src = reinterpret_cast<uintptr_t>(base + (size_t)row * A.row_stride);
                        const uint32_t pos = (uint32_t)(src & 15) + (uint32_t)(A.channels * 16 * sidx);   // byte in the ring row
                        const uint8_t *rp = ring + (g & 1) * RING_BUF_BYTES + r * RING_PITCH + (pos & ~15u);
                        const uint32_t ws = (pos & 15u) >> 2, bs = (pos & 3u) * 8;
                        uint32_t e[17];
#pragma unroll
                        for (int cidx = 0; cidx < 4; ++cidx) {
                            uint4 q = make_uint4(0, 0, 0, 0);
                            if (cidx < (A.channels == 3 ? 4 : 2) && (int)(pos >> 4) + cidx < RING_CHUNKS)
                                q = *reinterpret_cast<const uint4 *>(rp + 16 * cidx);
                            e[4 * cidx] = q.x; e[4 * cidx + 1] = q.y; e[4 * cidx + 2] = q.z; e[4 * cidx + 3] = q.w;
                        }
                        e[16] = 0;
                        uint32_t v[12];
                        switch (ws) {
                        case 0:
#pragma unroll
                            for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j], e[j + 1], bs);
                            break;
                        case 1:
#pragma unroll
                            for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 1], e[j + 2], bs);
                            break;
                        case 2:
#pragma unroll
                            for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 2], e[j + 3], bs);
                            break;
                        default:
#pragma unroll
                            for (int j = 0; j < 12; ++j) v[j] = __funnelshift_r(e[j + 3], e[j + 4], bs);
                            break;
                        }
                        out = A.channels == 1 ? make_uint4(v[0], v[1], v[2], v[3]) : lightness16(v);
                    }
                    stg_keep(reinterpret_cast<uint4 *>(dst + sidx * STRIP_BYTES + row * 16), out, pol_keep);
                }
            }
            worker_sync();                                   // the buffer may be refilled
        }
    }
}

}  // namespace

template <bool WRITE_MAP>
__global__ void __launch_bounds__(TC_THREADS, 1)
melf_match_fused_kernel(const MelfMatchTcArgs A)
{
    uint8_t *smem = melf_tc_smem;
    CtrlF *ctrl = reinterpret_cast<CtrlF *>(smem + F_SMEM_CTRL);
    uint32_t *Htab = reinterpret_cast<uint32_t *>(smem + F_SMEM_H);
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);          // provably warp-uniform
    const uint32_t slab_addr = smem_u32(smem + SMEM_SLABS), b_addr = smem_u32(smem + F_SMEM_B);
    uint8_t *scratch = A.slabs + (size_t)blockIdx.x * F_SCRATCH_PER_CTA;        // [slab 0][slab 1][S 0][S 1]
    uint32_t *wsum = reinterpret_cast<uint32_t *>(scratch + 2 * (size_t)SLAB_BYTES);

    if (tid == 0) {
        for (int s = 0; s < F_NSTAGE; ++s) {
            mbar_init(smem_u32(&ctrl->full[s]), 1);
            mbar_init(smem_u32(&ctrl->empty[s]), 1);
        }
        mbar_init(smem_u32(&ctrl->accum), 1);
        mbar_init(smem_u32(&ctrl->slab_full), 1);
        mbar_init(smem_u32(&ctrl->prep_done), F_WORKERS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     :: "r"(smem_u32(&ctrl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctrl->tmem_base;

    const long long n_pass = (A.n_items + 1) / 2;
    const int n_ksteps = 8 * A.n_pairs;                               // K-steps of 32 per item
    const int n_stages = (n_ksteps + F_KSTEPS - 1) / F_KSTEPS;        // per pass
    const int n_pre = n_stages < F_NSTAGE ? n_stages : F_NSTAGE;      // stages prefetched ahead of a pass
    const bool producer = warp == 0 && elect_one();
    const u64 pol_keep = policy_evict_last();

    // producer helpers.  Stage g of the running stage counter uses ring slot g % F_NSTAGE; its
    // "empty" barrier has completed (g / F_NSTAGE) phases when it may be refilled.
    auto load_stage = [&](uint32_t g, int s) {
        const uint32_t st = g % F_NSTAGE, ph = (g / F_NSTAGE) & 1;
        mbar_wait(smem_u32(&ctrl->empty[st]), ph ^ 1);
        const uint32_t bytes = (uint32_t)min(F_KSTEPS, n_ksteps - s * F_KSTEPS) * (2 * BSTRIP_BYTES);
        mbar_expect_tx(smem_u32(&ctrl->full[st]), bytes);
        bulk_load_keep(b_addr + st * F_STAGE_BYTES, A.toeplitz + (size_t)s * F_STAGE_BYTES, bytes,
                       smem_u32(&ctrl->full[st]), pol_keep);
    };
    auto load_slabs = [&](long long pass) {
        const int n = (2 * pass + 1 < A.n_items) ? 2 : 1;
        fence_async_all();                 // the workers' generic-proxy writes (global) / reads (slabs) first
        mbar_expect_tx(smem_u32(&ctrl->slab_full), (uint32_t)(n * SLAB_BYTES));
        for (int k = 0; k < n; ++k)
            bulk_load_keep(slab_addr + k * SLAB_BYTES, scratch + (size_t)k * SLAB_BYTES, SLAB_BYTES,
                           smem_u32(&ctrl->slab_full), pol_keep);
    };
    auto pass_items = [&](long long pass, Item (&its)[2]) {
        its[0] = get_item(A, 2 * pass);
        its[1] = get_item(A, 2 * pass + 1);
    };

    uint32_t stage_count = 0;       // Toeplitz stages consumed before this pass
    uint32_t pass_count = 0;

    // ---- prologue: slabs of this CTA's first pass --------------------------------------------
    if ((long long)blockIdx.x < n_pass) {
        if (warp >= 2) {
            Item its[2];
            pass_items(blockIdx.x, its);
            stage_slabs(A, its, scratch, warp - 2, lane);
            fence_async_all();
            mbar_arrive(smem_u32(&ctrl->prep_done));
        } else if (producer) {
            for (int s = 0; s < n_pre; ++s) load_stage((uint32_t)s, s);
            mbar_wait(smem_u32(&ctrl->prep_done), 0);
            load_slabs(blockIdx.x);
        }
    }
    __syncwarp();

    for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x, ++pass_count) {
        Item items[2];
        pass_items(pass, items);
        const bool has_next = pass + gridDim.x < n_pass;
        long long *tl = (A.timeline && blockIdx.x == 0 && pass_count < 8) ? A.timeline + pass_count * 8 : nullptr;

        if (warp == 0) {
            // ---- producer: the rest of the Toeplitz stream --------------------------------------
            if (producer)
                for (int s = n_pre; s < n_stages; ++s) load_stage(stage_count + (uint32_t)s, s);
        } else if (warp == 1) {
            // ---- issuer: the whole warp walks the loop, one elected lane issues -----------------
            const bool lead = elect_one();
            if (tl && lead) tl[0] = clock64();
            mbar_wait(smem_u32(&ctrl->slab_full), pass_count & 1);
            tc_fence_after();
            if (tl && lead) tl[1] = clock64();
            const uint32_t idesc0 = make_idesc((uint32_t)items[0].n_cols);
            const uint32_t idesc1 = make_idesc((uint32_t)items[1].n_cols);
            uint32_t g = stage_count;
            long long waited = 0;                   // debug: clocks spent waiting for Toeplitz stages
            for (int s = 0; s < n_stages; ++s, ++g) {
                const uint32_t st = g % F_NSTAGE, ph = (g / F_NSTAGE) & 1;
                const long long w0 = tl ? clock64() : 0;
                mbar_wait(smem_u32(&ctrl->full[st]), ph);
                tc_fence_after();
                if (tl) waited += clock64() - w0;
                if (lead) {
                    const int kk0 = s * F_KSTEPS;                                     // first K-step of the stage
                    const uint32_t abase = b_addr + st * F_STAGE_BYTES;               // Toeplitz K-steps = A
#pragma unroll
                    for (int slot = 0; slot < 2; ++slot) {
                        if (!(slot ? items[1].valid : items[0].valid)) continue;
                        const uint32_t d = tmem + (uint32_t)(slot * NCOLS);
#pragma unroll
                        for (int j = 0; j < F_KSTEPS; ++j) {
                            const int kk = kk0 + j;
                            if (kk < n_ksteps) {
                                // template rows 2p, 2p+1 against slab rows 2p.. (= B), K-step ks of the pair
                                const int pp = kk >> 3, ks = kk & 7;
                                const uint32_t bbase = slab_addr + slot * SLAB_BYTES + (2 * pp) * 16 + ks * 2 * STRIP_BYTES;
                                umma_i8(d, umma_desc(abase + j * 2 * BSTRIP_BYTES, BSTRIP_BYTES),
                                        umma_desc(bbase, STRIP_BYTES),
                                        slot ? idesc1 : idesc0, kk > 0 ? 1u : 0u);
                            }
                        }
                    }
                    umma_commit(smem_u32(&ctrl->empty[st]));
                }
                __syncwarp();
            }
            if (lead) umma_commit(smem_u32(&ctrl->accum));
            if (tl && lead) tl[2] = tl[0] + waited;   // reported relative to tl[0] like the stamps
        } else {
            // ---- workers: window sums of this pass, slabs of the next one -------------------------
            const int wt = tid - 64;
            mbar_wait(smem_u32(&ctrl->slab_full), pass_count & 1);
#pragma unroll 1
            for (int slot = 0; slot < 2; ++slot) {
                const Item cur = slot ? items[1] : items[0];
#ifdef MELF_TC_EXP
                if (!(MELF_TC_EXP & 1))
#endif
                if (cur.valid)
                    window_sums(A, cur, smem + SMEM_SLABS + slot * SLAB_BYTES, Htab, wsum + slot * WSUM_WORDS, wt);
            }
            if (tl && wt == 0) tl[6] = clock64();
            if (has_next) {
                Item nxt[2];
                pass_items(pass + gridDim.x, nxt);
#if defined(MELF_TC_EXP) && (MELF_TC_EXP & 4)
                // experiment: keep the workers off the issuer's scheduler (warp % 4 == 1)
                if ((warp & 3) != 1) stage_slabs(A, nxt, scratch, warp - 2 - ((warp - 2) >> 2), lane, 11);
#elif defined(MELF_TC_EXP) && (MELF_TC_EXP & 8)
                // experiment: keep the workers off the producer's scheduler (warp % 4 == 0)
                if ((warp & 3) != 0) stage_slabs(A, nxt, scratch, warp - 2 - ((warp - 1) >> 2), lane, 11);
#else
#ifdef MELF_TC_EXP
                if (!(MELF_TC_EXP & 2))
#endif
                stage_slabs(A, nxt, scratch, warp - 2, lane);
#endif
            }
            if (tl && wt == 0) tl[7] = clock64();
            fence_async_all();
            mbar_arrive(smem_u32(&ctrl->prep_done));
        }
        stage_count += (uint32_t)n_stages;
        __syncwarp();

        // ---- epilogue ----------------------------------------------------------------------------
        // warp -> TMEM lane quadrant qd (hardware rule: warp % 4), item slot, half of the columns.
        // Lane pair (2j, 2j+1) holds D[2x][.] and D[2x+1][.] of result column x = 16 qd + j; of
        // every 16-row chunk the even lane scores rows 0..7 and the odd lane rows 8..15.
        const int qd = warp & 3, slot = (warp >> 2) & 1, half = warp >> 3;
        const Item it = slot ? items[1] : items[0];
        const int n_chunks = it.n_cols >> 4;
        const int c_mid = (n_chunks + 1) >> 1;
        const int c0 = half ? c_mid : 0, c1 = half ? n_chunks : c_mid;      // <= 5 chunks
        const int x = 16 * qd + (lane >> 1);                       // result column inside the item
        const int odd = lane & 1;
        const bool col_ok = it.valid && x < it.w_c;

        mbar_wait(smem_u32(&ctrl->prep_done), (pass_count + 1) & 1);       // window sums + next slabs written
        const uint32_t *S = wsum + slot * WSUM_WORDS + x;
        uint32_t sw[5][8];
#pragma unroll
        for (int k = 0; k < 5; ++k)
            if (c0 + k < c1 && col_ok) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int yl = 16 * (c0 + k) + 8 * odd + j;
                    sw[k][j] = yl < it.h_c ? __ldcg(S + yl * 64) : 0u;
                }
            }

        mbar_wait(smem_u32(&ctrl->accum), pass_count & 1);
        tc_fence_after();
        if (tl && tid == 64) tl[3] = clock64();
        // every MMA of this pass has completed: slabs and Toeplitz ring are free again
        if (producer && has_next) {
            load_slabs(pass + gridDim.x);
            for (int s = 0; s < n_pre; ++s) load_stage(stage_count + (uint32_t)s, s);
        }
        __syncwarp();

        const uint32_t taddr = tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(slot * NCOLS);
        const uint32_t n_t32 = (uint32_t)A.n_t, sum_t32 = (uint32_t)A.sum_t;      // <= 193 * 128 (* 255)
        // Running arg-max per lane.  The lane visits its raster indices in increasing order, so a
        // candidate replaces the running best only if its float32 score is strictly greater.  Scores
        // are a monotone function of the numerator and more than 4 float32 ulps of numerator distance
        // always separate the rounded values: beyond `far` (that distance for the largest possible
        // |numerator| = n_t * 255 * sum_t) the integer comparison decides, closer calls (rare) round.
        const long long kNone = -(1ll << 62);
        const long long far = (((long long)((u64)n_t32 * 255ull) * (long long)sum_t32 + A.n_t) >> 21) + 1;
        long long best_num = kNone;
        uint32_t best_idx = 0;
        float *map = WRITE_MAP ? A.result + (size_t)it.frame * A.geom.res_w * A.geom.res_h : nullptr;
        if (it.valid) {
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const int c = c0 + k;
                if (c < c1) {                                // warp-uniform
                    uint32_t d16[16], d[17];
                    tmem_ld16(taddr + 16 * c, d16);
                    tmem_ld1(taddr + 16 * c + 16, d[16]);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) d[i] = d16[i];
                    const int yl0 = 16 * c + 8 * odd;
                    uint32_t idx = (uint32_t)(it.y0 + yl0) * (uint32_t)A.geom.res_w + (uint32_t)(it.x0 + x);
#pragma unroll
                    for (int j = 0; j < 8; ++j, idx += (uint32_t)A.geom.res_w) {
                        // even sends D[2x][y0+8+j] and receives D[2x+1][y0+j+1]; odd the other way round
                        const uint32_t other = __shfl_xor_sync(0xffffffffu, odd ? d[j + 1] : d[8 + j], 1);
                        const uint32_t mine = odd ? d[9 + j] : d[j];
                        // all four factors are below 2^32: two 32 x 32 -> 64-bit multiplies
                        const long long num = (long long)((u64)n_t32 * (u64)(mine + other)) -
                                              (long long)((u64)sum_t32 * (u64)sw[k][j]);
                        const bool ok = col_ok && yl0 + j < it.h_c;
                        if (WRITE_MAP) { if (ok) map[idx] = score_of(num, A.n_t); }
                        const long long gain = num - best_num;
                        bool take = ok && gain > far;
                        if (ok && gain > 0 && !take)         // close call
                            take = score_of(num, A.n_t) > score_of(best_num, A.n_t);
                        best_num = take ? num : best_num;
                        best_idx = take ? idx : best_idx;
                    }
                }
            }
        }
        const bool has = best_num != kNone;
        if (tl && tid == 64) tl[4] = clock64();
        u64 best = has ? melf_score_key(score_of(best_num, A.n_t), best_idx) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if (lane == 0) ctrl->red[pass_count & 1][warp] = best;
        tc_fence_before();
        __syncthreads();            // TMEM fully read before the next pass's first MMA overwrites it
        tc_fence_after();
        if (tl && tid == 64) tl[5] = clock64();
        if (tid < 2) {
            // warps of slot s: 4 s .. 4 s + 3 and 8 + 4 s .. 8 + 4 s + 3
            const Item mine = tid ? items[1] : items[0];
            u64 b = 0;
            for (int w8 = 0; w8 < 8; ++w8) {
                const u64 k = ctrl->red[pass_count & 1][(w8 & 3) + 4 * tid + 8 * (w8 >> 2)];
                b = k > b ? k : b;
            }
            if (mine.valid && b) atomicMax(A.match_key + mine.frame, b);
        }
        __syncwarp();               // red[] alternates per pass: no thread can be two passes behind
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

size_t melf_match_fused_scratch_bytes(int num_sms) { return (size_t)num_sms * F_SCRATCH_PER_CTA; }

cudaError_t melf_launch_match_fused(const MelfMatchTcArgs &args, int num_sms, cudaStream_t stream)
{
    cudaError_t e = cudaFuncSetAttribute(melf_match_fused_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_TOTAL);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(melf_match_fused_kernel<true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_TOTAL);
    if (e != cudaSuccess) return e;
    if (args.n_items <= 0) return cudaSuccess;
    const long long n_pass = (args.n_items + 1) / 2;
    const unsigned grid = (unsigned)(n_pass < num_sms ? n_pass : num_sms);
    if (args.result)
        melf_match_fused_kernel<true><<<grid, TC_THREADS, F_SMEM_TOTAL, stream>>>(args);
    else
        melf_match_fused_kernel<false><<<grid, TC_THREADS, F_SMEM_TOTAL, stream>>>(args);
    return cudaGetLastError();
}


// =======================================================================================
// Overlapped variant of the fused kernel: the drain of pass k runs under the MMAs of pass k+1.
// Three accumulator buffers rotate through tensor memory (item i of the CTA -> buffer i % 3,
// 432 of 512 columns); warps 0 (producer) and 1 (issuer) never drain, so the issuer walks straight
// into the next pass; the 14 worker warps drain (slot 0 first: its buffer is the one pass k+1 needs
// for its second item), then do their window sums / slab staging.  No block-wide barrier inside
// the pass loop: every hand-over is an mbarrier.
// =======================================================================================
namespace {

struct CtrlO {
    unsigned long long full[F_NSTAGE];
    unsigned long long empty[F_NSTAGE];
    unsigned long long accum;
    unsigned long long slab_full;
    unsigned long long prep_done;
    unsigned long long drained[2];     // accumulator of the pass's slot 0 / slot 1 read out (14 warp arrivals)
    unsigned long long stage[F_WORKER_WARPS];   // per worker warp: its staging buffer has been filled (bulk copies)
    unsigned long long dummy;          // MELF_TC_EXPERIMENT only
    uint32_t tmem_base;
    uint32_t pad;
};

}  // namespace

template <bool WRITE_MAP>
__global__ void __launch_bounds__(TC_THREADS, 1)
melf_match_fused_ov_kernel(const MelfMatchTcArgs A, const __grid_constant__ CUtensorMap frames_map)
{
    uint8_t *smem = melf_tc_smem;
    CtrlO *ctrl = reinterpret_cast<CtrlO *>(smem + F_SMEM_CTRL);
    uint32_t *Htab = reinterpret_cast<uint32_t *>(smem + F_SMEM_H);
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);          // provably warp-uniform
    const uint32_t slab_addr = smem_u32(smem + SMEM_SLABS), b_addr = smem_u32(smem + F_SMEM_B);
    // per-CTA scratch: [slab 0][slab 1][S 0][S 1] of even passes, then [S 0][S 1] of odd passes
    uint8_t *scratch = A.slabs + (size_t)blockIdx.x * F_SCRATCH_PER_CTA_OV;
    uint32_t *wsum_base = reinterpret_cast<uint32_t *>(scratch + 2 * (size_t)SLAB_BYTES);

    if (tid == 0) {
        for (int s = 0; s < F_NSTAGE; ++s) {
            mbar_init(smem_u32(&ctrl->full[s]), 1);
            mbar_init(smem_u32(&ctrl->empty[s]), 1);
        }
        mbar_init(smem_u32(&ctrl->accum), 1);
        mbar_init(smem_u32(&ctrl->slab_full), 1);
        mbar_init(smem_u32(&ctrl->prep_done), F_WORKERS);
        mbar_init(smem_u32(&ctrl->drained[0]), F_WORKER_WARPS);
        mbar_init(smem_u32(&ctrl->drained[1]), F_WORKER_WARPS);
        for (int w = 0; w < F_WORKER_WARPS; ++w) mbar_init(smem_u32(&ctrl->stage[w]), 1);
        mbar_init(smem_u32(&ctrl->dummy), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     :: "r"(smem_u32(&ctrl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctrl->tmem_base;

    const long long n_pass = (A.n_items + 1) / 2;
    const int n_ksteps = 8 * A.n_pairs;                               // K-steps of 32 per item
    const int n_stages = (n_ksteps + F_KSTEPS - 1) / F_KSTEPS;        // per pass
    const int n_pre = n_stages < F_NSTAGE ? n_stages : F_NSTAGE;      // stages prefetched ahead of a pass
    // K-steps of a pair that can hold a template byte: row n of the Toeplitz band ends at k = n + tmpl_w - 1 <= 62 +
    // tmpl_w, so the K-steps from ceil((63 + tmpl_w) / 32) on are all zero (the eighth one for templates up to 161
    // wide) and are neither loaded nor multiplied.  (A stage is one pair only with 8 K-steps per stage.)
    const int kpp_stage = F_KSTEPS == 8 ? A.kpp : F_KSTEPS;
    const u64 pol_keep = policy_evict_last();
    auto pass_items = [&](long long pass, Item (&its)[2]) {
        its[0] = get_item(A, 2 * pass);
        its[1] = get_item(A, 2 * pass + 1);
    };
    // accumulator buffer of slot `slot` of this CTA's pc-th pass
    auto acc_col = [&](uint32_t pc, int slot) -> uint32_t { return ((2u * pc + (uint32_t)slot) % 3u) * (uint32_t)NCOLS; };

    if ((long long)blockIdx.x < n_pass) {
    if (warp == 0) {
        // =============================== producer ===============================================
        if (elect_one()) {
            auto load_stage = [&](uint32_t g, int s) {
                const uint32_t st = g % F_NSTAGE, ph = (g / F_NSTAGE) & 1;
                mbar_wait(smem_u32(&ctrl->empty[st]), ph ^ 1);
                const uint32_t bytes = (uint32_t)min(kpp_stage, n_ksteps - s * F_KSTEPS) * (2 * BSTRIP_BYTES);
                mbar_expect_tx(smem_u32(&ctrl->full[st]), bytes);
                bulk_load_keep(b_addr + st * F_STAGE_BYTES, A.toeplitz + (size_t)s * F_STAGE_BYTES, bytes,
                               smem_u32(&ctrl->full[st]), pol_keep);
            };
            auto load_slabs = [&](long long pass) {
                const int n = (2 * pass + 1 < A.n_items) ? 2 : 1;
                fence_async_all();         // the workers' generic-proxy writes (global) / reads (slabs) first
                mbar_expect_tx(smem_u32(&ctrl->slab_full), (uint32_t)(n * SLAB_BYTES));
                for (int k = 0; k < n; ++k)
                    bulk_load_keep(slab_addr + k * SLAB_BYTES, scratch + (size_t)k * SLAB_BYTES, SLAB_BYTES,
                                   smem_u32(&ctrl->slab_full), pol_keep);
            };
            uint32_t stage_count = 0, pc = 0;
            for (int s = 0; s < n_pre; ++s) load_stage((uint32_t)s, s);
            mbar_wait(smem_u32(&ctrl->prep_done), 0);
            load_slabs(blockIdx.x);
            for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x, ++pc) {
                for (int s = n_pre; s < n_stages; ++s) load_stage(stage_count + (uint32_t)s, s);
                stage_count += (uint32_t)n_stages;
                // every MMA of this pass complete (slabs, ring free) and the next slabs staged
                mbar_wait(smem_u32(&ctrl->accum), pc & 1);
                mbar_wait(smem_u32(&ctrl->prep_done), (pc + 1) & 1);
                if (pass + gridDim.x < n_pass) {
                    load_slabs(pass + gridDim.x);
                    for (int s = 0; s < n_pre; ++s) load_stage(stage_count + (uint32_t)s, s);
                }
            }
        }
    } else if (warp == 1) {
        // =============================== issuer =================================================
        // the whole warp walks the loop (every lane polls the barriers), one elected lane issues
        const bool lead = elect_one();
        uint32_t g = 0, pc = 0;
        for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x, ++pc) {
            Item items[2];
            pass_items(pass, items);
            long long *tl = (A.timeline && blockIdx.x == 0 && pc < 8) ? A.timeline + pc * 8 : nullptr;
            if (tl && lead) tl[0] = clock64();
            mbar_wait(smem_u32(&ctrl->slab_full), pc & 1);
            tc_fence_after();
            if (tl && lead) tl[1] = clock64();
            const uint32_t idesc0 = make_idesc((uint32_t)items[0].n_cols);
            const uint32_t idesc1 = make_idesc((uint32_t)items[1].n_cols);
            const uint32_t d0 = tmem + acc_col(pc, 0), d1 = tmem + acc_col(pc, 1);
            long long waited = 0;
            for (int s = 0; s < n_stages; ++s, ++g) {
                const uint32_t st = g % F_NSTAGE, ph = (g / F_NSTAGE) & 1;
                const long long w0 = tl ? clock64() : 0;
                mbar_wait(smem_u32(&ctrl->full[st]), ph);
                tc_fence_after();
                if (tl) waited += clock64() - w0;
                const int kk0 = s * F_KSTEPS;                                     // first K-step of the stage
                const uint32_t abase = b_addr + st * F_STAGE_BYTES;               // Toeplitz K-steps = A
                if (lead && items[0].valid && !(A.debug_skip & 8)) {
#pragma unroll
                    for (int j = 0; j < F_KSTEPS; ++j) {
                        const int kk = kk0 + j;
                        if (kk < n_ksteps && j < kpp_stage) {
                            const int pp = kk >> 3, ks = kk & 7;
                            umma_i8(d0, umma_desc(abase + j * 2 * BSTRIP_BYTES, BSTRIP_BYTES),
                                    umma_desc(slab_addr + (2 * pp) * 16 + ks * 2 * STRIP_BYTES, STRIP_BYTES),
                                    idesc0, kk > 0 ? 1u : 0u);
                        }
                    }
                }
#if MELF_TC_EXPERIMENT == 1
                if (lead) umma_commit(smem_u32(&ctrl->dummy));      // what does a commit cost the MMA stream?
#elif MELF_TC_EXPERIMENT == 2
                tc_fence_after();
#elif MELF_TC_EXPERIMENT == 3
                mbar_wait(smem_u32(&ctrl->full[st]), ph);
                __syncwarp();
#endif
                if (s == 0 && pc >= 1) {
                    // slot 1's buffer held slot 0 of the previous pass: wait until it has been read out
                    mbar_wait(smem_u32(&ctrl->drained[0]), (pc - 1) & 1);
                    tc_fence_after();
                }
                if (lead) {
                    if (items[1].valid && !(A.debug_skip & 8)) {
#pragma unroll
                        for (int j = 0; j < F_KSTEPS; ++j) {
                            const int kk = kk0 + j;
                            if (kk < n_ksteps && j < kpp_stage) {
                                const int pp = kk >> 3, ks = kk & 7;
                                umma_i8(d1, umma_desc(abase + j * 2 * BSTRIP_BYTES, BSTRIP_BYTES),
                                        umma_desc(slab_addr + SLAB_BYTES + (2 * pp) * 16 + ks * 2 * STRIP_BYTES, STRIP_BYTES),
                                        idesc1, kk > 0 ? 1u : 0u);
                            }
                        }
                    }
                    umma_commit(smem_u32(&ctrl->empty[st]));
                }
                __syncwarp();
            }
            // slot 0 of the NEXT pass reuses the buffer of slot 1 of the PREVIOUS pass: it must have been
            // read out (waiting here, before this pass's commit, keeps the barrier at most one phase ahead)
            if (pc >= 1) {
                mbar_wait(smem_u32(&ctrl->drained[1]), (pc - 1) & 1);
                tc_fence_after();
            }
            if (lead) umma_commit(smem_u32(&ctrl->accum));
            if (tl && lead) { tl[2] = tl[0] + waited; tl[3] = clock64(); }
        }
    } else {
        // =============================== workers ================================================
        const int wt = tid - 64, wwarp = warp - 2;
        const int qd = warp & 3;                               // TMEM lane quadrant this warp can read
        // drain duty inside the quadrant: quadrants 0, 1 have the warps 4,8,12 / 5,9,13; 2, 3 have four
        const int q_first = qd < 2 ? 4 + qd : qd;
        const int q_members = (TC_THREADS / 32 - 1 - q_first) / 4 + 1;
        const int q_index = (warp - q_first) >> 2;
        const int x = 16 * qd + (lane >> 1);                       // result column inside the item
        const int odd = lane & 1;
        const uint32_t n_t32 = (uint32_t)A.n_t, sum_t32 = (uint32_t)A.sum_t;      // <= 193 * 128 (* 255)
        const long long kNone = -(1ll << 62);
        const long long far = (((long long)((u64)n_t32 * 255ull) * (long long)sum_t32 + A.n_t) >> 21) + 1;
        uint32_t stage_phase = 0;                              // parity of this warp's staging barrier
        (void)stage_phase;
        {
            // prologue: slabs of this CTA's first pass
            Item its[2];
            pass_items(blockIdx.x, its);
#if MELF_STAGE_RING
            stage_slabs_ring(A, its, scratch, smem + F_SMEM_H, wt);
#elif MELF_STAGE_TMA
            if (A.use_tmap)
                stage_slabs_tmap(A, &frames_map, its, scratch, smem + F_SMEM_H + wwarp * CO_BUF_BYTES, smem_u32(&ctrl->stage[wwarp]), stage_phase, wwarp, lane);
            else
                stage_slabs_tma(A, its, scratch, smem + F_SMEM_H + wwarp * CO_BUF_BYTES, smem_u32(&ctrl->stage[wwarp]), stage_phase, wwarp, lane);
#elif MELF_STAGE_COALESCED
            stage_slabs_co(A, its, scratch, smem + F_SMEM_H + wwarp * CO_BUF_BYTES, wwarp, lane);
#else
            stage_slabs(A, its, scratch, wwarp, lane);
#endif
            fence_async_all();
            mbar_arrive(smem_u32(&ctrl->prep_done));
        }
        uint32_t pc = 0;
        for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x, ++pc) {
            Item items[2];
            pass_items(pass, items);
            const bool has_next = pass + gridDim.x < n_pass;
            long long *tl = (A.timeline && blockIdx.x == 0 && pc < 8) ? A.timeline + pc * 8 : nullptr;
            uint32_t *wsum = wsum_base + (pc & 1) * 2 * WSUM_WORDS;

            // ---- window sums of this pass, slabs of the next one ----------------------------------
#if MELF_STAGE_L2_PREFETCH
            if (has_next) {
                Item nxt[2];
                pass_items(pass + gridDim.x, nxt);
                prefetch_items_l2(A, nxt, wt);
            }
#endif
            mbar_wait(smem_u32(&ctrl->slab_full), pc & 1);
#pragma unroll 1
            for (int slot = 0; slot < 2; ++slot) {
                const Item cur = slot ? items[1] : items[0];
                if (cur.valid && !(A.debug_skip & 2))
                    window_sums(A, cur, smem + SMEM_SLABS + slot * SLAB_BYTES, Htab, wsum + slot * WSUM_WORDS, wt);
            }
            if (tl && wt == 0) tl[6] = clock64();
            if (has_next && !(A.debug_skip & 1)) {
                Item nxt[2];
                pass_items(pass + gridDim.x, nxt);
#if MELF_STAGE_RING
                stage_slabs_ring(A, nxt, scratch, smem + F_SMEM_H, wt);
#elif MELF_STAGE_TMA
                if (A.use_tmap)
                    stage_slabs_tmap(A, &frames_map, nxt, scratch, smem + F_SMEM_H + wwarp * CO_BUF_BYTES, smem_u32(&ctrl->stage[wwarp]), stage_phase, wwarp, lane);
                else
                    stage_slabs_tma(A, nxt, scratch, smem + F_SMEM_H + wwarp * CO_BUF_BYTES, smem_u32(&ctrl->stage[wwarp]), stage_phase, wwarp, lane);
#elif MELF_STAGE_COALESCED
                stage_slabs_co(A, nxt, scratch, smem + F_SMEM_H + wwarp * CO_BUF_BYTES, wwarp, lane);
#else
                stage_slabs(A, nxt, scratch, wwarp, lane);
#endif
            }
            if (tl && wt == 0) tl[7] = clock64();
            fence_async_all();
            mbar_arrive(smem_u32(&ctrl->prep_done));

            // ---- drain (under the MMAs of the next pass) ------------------------------------------
            mbar_wait(smem_u32(&ctrl->prep_done), (pc + 1) & 1);   // every worker's window sums are visible
            mbar_wait(smem_u32(&ctrl->accum), pc & 1);
            tc_fence_after();
#pragma unroll 1
            for (int slot = 0; slot < 2; ++slot) {
                const Item it = slot ? items[1] : items[0];
                const int n_chunks = it.n_cols >> 4;
                const bool col_ok = it.valid && x < it.w_c;
                const uint32_t taddr = tmem + ((uint32_t)(32 * qd) << 16) + acc_col(pc, slot);
                const uint32_t *S = wsum + slot * WSUM_WORDS + x;
                float *map = WRITE_MAP ? A.result + (size_t)it.frame * A.geom.res_w * A.geom.res_h : nullptr;
                long long best_num = kNone;
                uint32_t best_idx = 0;
                if (A.debug_skip & 4) {
                    // measurement only: accumulators are not read
                } else
#if MELF_DRAIN_WIDE
                if (it.valid) {
                    // 32 result rows per step: one x32 (+ x1) accumulator load instead of two x16 (+ x1) -- every
                    // tcgen05.ld queues behind the MMAs in flight, so fewer, larger loads wait less often; the
                    // even lane of a pair scores rows +0..15, the odd lane rows +16..31
                    const int n_wide = (it.n_cols + 31) >> 5;
                    uint32_t sw_nx[16];
                    auto fetch_s = [&](int c) {
                        const int y0c = 32 * c + 16 * odd;
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            sw_nx[j] = (c < n_wide && col_ok && y0c + j < it.h_c) ? __ldcg(S + (y0c + j) * 64) : 0u;
                    };
                    fetch_s(q_index);
#pragma unroll 1
                    for (int c = q_index; c < n_wide; c += q_members) {
                        const int yl0 = 32 * c + 16 * odd;
                        uint32_t d[33], sw[16];
                        tmem_ld32(taddr + 32 * c, d);
                        tmem_ld1(taddr + 32 * c + 32, d[32]);
#pragma unroll
                        for (int j = 0; j < 16; ++j) sw[j] = sw_nx[j];
                        fetch_s(c + q_members);
                        tmem_ld_wait();
                        uint32_t idx = (uint32_t)(it.y0 + yl0) * (uint32_t)A.geom.res_w + (uint32_t)(it.x0 + x);
#pragma unroll
                        for (int j = 0; j < 16; ++j, idx += (uint32_t)A.geom.res_w) {
                            // even sends D[2x][y0+16+j] and receives D[2x+1][y0+j+1]; odd the other way round
                            const uint32_t other = __shfl_xor_sync(0xffffffffu, odd ? d[j + 1] : d[16 + j], 1);
                            const uint32_t mine = odd ? d[17 + j] : d[j];
                            const long long num = (long long)((u64)n_t32 * (u64)(mine + other)) -
                                                  (long long)((u64)sum_t32 * (u64)sw[j]);
                            const bool ok = col_ok && yl0 + j < it.h_c;
                            if (WRITE_MAP) { if (ok) map[idx] = score_of(num, A.n_t); }
                            const long long gain = num - best_num;
                            bool take = ok && gain > far;
                            if (ok && gain > 0 && !take)
                                take = score_of(num, A.n_t) > score_of(best_num, A.n_t);
                            best_num = take ? num : best_num;
                            best_idx = take ? idx : best_idx;
                        }
                    }
                }
#else
                if (it.valid) {
                    // the window sums of a chunk are fetched (L2) while the previous chunk is scored
                    uint32_t sw_nx[8];
                    auto fetch_s = [&](int c) {
                        const int y0c = 16 * c + 8 * odd;
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            sw_nx[j] = (c < n_chunks && col_ok && y0c + j < it.h_c) ? __ldcg(S + (y0c + j) * 64) : 0u;
                    };
                    fetch_s(q_index);
                    // ... and so are its accumulator columns (a tcgen05.ld queues behind the MMAs in flight)
                    uint32_t d16_nx[16], d1_nx = 0;
                    if (q_index < n_chunks) {
                        tmem_ld16(taddr + 16 * q_index, d16_nx);
                        tmem_ld1(taddr + 16 * q_index + 16, d1_nx);
                    }
#pragma unroll 1
                    for (int c = q_index; c < n_chunks; c += q_members) {
                        const int yl0 = 16 * c + 8 * odd;
                        uint32_t sw[8], d[17];
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 8; ++j) sw[j] = sw_nx[j];
#pragma unroll
                        for (int i = 0; i < 16; ++i) d[i] = d16_nx[i];
                        d[16] = d1_nx;
                        fetch_s(c + q_members);
                        if (c + q_members < n_chunks) {
                            tmem_ld16(taddr + 16 * (c + q_members), d16_nx);
                            tmem_ld1(taddr + 16 * (c + q_members) + 16, d1_nx);
                        }
                        uint32_t idx = (uint32_t)(it.y0 + yl0) * (uint32_t)A.geom.res_w + (uint32_t)(it.x0 + x);
#pragma unroll
                        for (int j = 0; j < 8; ++j, idx += (uint32_t)A.geom.res_w) {
                            // even sends D[2x][y0+8+j] and receives D[2x+1][y0+j+1]; odd the other way round
                            const uint32_t other = __shfl_xor_sync(0xffffffffu, odd ? d[j + 1] : d[8 + j], 1);
                            const uint32_t mine = odd ? d[9 + j] : d[j];
                            const long long num = (long long)((u64)n_t32 * (u64)(mine + other)) -
                                                  (long long)((u64)sum_t32 * (u64)sw[j]);
                            const bool ok = col_ok && yl0 + j < it.h_c;
                            if (WRITE_MAP) { if (ok) map[idx] = score_of(num, A.n_t); }
                            // see melf_match_fused_kernel: raster order per lane, rounding only for close calls
                            const long long gain = num - best_num;
                            bool take = ok && gain > far;
                            if (ok && gain > 0 && !take)
                                take = score_of(num, A.n_t) > score_of(best_num, A.n_t);
                            best_num = take ? num : best_num;
                            best_idx = take ? idx : best_idx;
                        }
                    }
                }
#endif
                // this warp is done with the slot's accumulator
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&ctrl->drained[slot]));
                u64 best = best_num != kNone ? melf_score_key(score_of(best_num, A.n_t), best_idx) : 0ull;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
                    best = other > best ? other : best;
                }
                if (lane == 0 && it.valid && best) atomicMax(A.match_key + it.frame, best);
            }
            if (tl && wt == 0) tl[4] = clock64();
        }
    }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

size_t melf_match_fused_ov_scratch_bytes(int num_sms) { return (size_t)num_sms * F_SCRATCH_PER_CTA_OV; }

// Tensor map of the caller's frame batch for stage_slabs_tmap: (bytes of a row, rows, frames), box = 8 rows x
// (64 pixels + 16 bytes).  False when the batch cannot be described (base or strides not multiples of 16,
// driver entry point missing): the kernel then stages with the per-row bulk copies, which take any alignment.
static bool make_frames_map(const MelfMatchTcArgs &a, CUtensorMap *map)
{
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                 const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = [] {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) fn = nullptr;
        return reinterpret_cast<EncodeFn>(fn);
    }();
    static const bool off = std::getenv("MELF_STAGE_TMAP") && !std::atoi(std::getenv("MELF_STAGE_TMAP"));
    if (!encode || off || a.img_w <= 0 || a.img_h <= 0 || a.img_n <= 0) return false;
    if ((reinterpret_cast<uintptr_t>(a.img) & 15) || (a.row_stride & 15) || (a.frame_stride & 15)) return false;
    if (a.channels != 1 && a.channels != 3) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)a.img_w * (cuuint64_t)a.channels, (cuuint64_t)a.img_h, (cuuint64_t)a.img_n};
    const cuuint64_t strides[2] = {(cuuint64_t)a.row_stride, (cuuint64_t)a.frame_stride};
    const cuuint32_t box[3] = {(cuuint32_t)(64 * a.channels + 16), 8, 1};
    const cuuint32_t elem[3] = {1, 1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t *>(a.img), dims, strides, box, elem,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

cudaError_t melf_launch_match_fused_ov(const MelfMatchTcArgs &args_in, int num_sms, cudaStream_t stream)
{
    MelfMatchTcArgs args = args_in;
    alignas(64) CUtensorMap frames_map;
    std::memset(&frames_map, 0, sizeof frames_map);
    args.use_tmap = make_frames_map(args, &frames_map) ? 1 : 0;
    cudaError_t e = cudaFuncSetAttribute(melf_match_fused_ov_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_TOTAL);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(melf_match_fused_ov_kernel<true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_TOTAL);
    if (e != cudaSuccess) return e;
    if (args.n_items <= 0) return cudaSuccess;
    const long long n_pass = (args.n_items + 1) / 2;
    const unsigned grid = (unsigned)(n_pass < num_sms ? n_pass : num_sms);
    if (args.result)
        melf_match_fused_ov_kernel<true><<<grid, TC_THREADS, F_SMEM_TOTAL, stream>>>(args, frames_map);
    else
        melf_match_fused_ov_kernel<false><<<grid, TC_THREADS, F_SMEM_TOTAL, stream>>>(args, frames_map);
    return cudaGetLastError();
}

// ---- host side ---------------------------------------------------------------------------

bool melf_match_tc_supported(int tmpl_w, int tmpl_h)
{
    return tmpl_w >= 1 && tmpl_w + 63 <= KS * 16 && tmpl_h >= 1 && tmpl_h <= 128;
}

size_t melf_match_tc_toeplitz_bytes(int tmpl_h)
{
    return (size_t)((tmpl_h + 1) / 2) * BSTAGE_BYTES;
}

// Toeplitz operand image: [pair p][strip s][m 0..127][16 bytes]; row m = 2 n + q belongs to
// result column n and template row 2p + q (zero beyond the template); byte b of strip s is
// K index k = 16 s + b and holds T(2p + q, k - n).
void melf_match_tc_build_toeplitz(const uint8_t *tmpl, int tmpl_w, int tmpl_h, uint8_t *out)
{
    const int n_pairs = (tmpl_h + 1) / 2;
    for (int p = 0; p < n_pairs; ++p)
        for (int s = 0; s < KS; ++s)
            for (int m = 0; m < BM; ++m)
                for (int b = 0; b < 16; ++b) {
                    const int row = 2 * p + (m & 1), j = 16 * s + b - (m >> 1);
                    uint8_t v = 0;
                    if (row < tmpl_h && j >= 0 && j < tmpl_w) v = tmpl[(size_t)row * tmpl_w + j];
                    out[(((size_t)p * KS + s) * BM + m) * 16 + b] = v;
                }
}

// fills the tiling fields of args; returns the number of passes
long long melf_match_tc_plan(MelfMatchTcArgs &a, int n_frames)
{
    const int max_h = SLAB_DATA_ROWS - a.tmpl_h;       // image rows r + i <= h_c + th - 2 <= 254
    const int cap = max_h < SROWS ? max_h : SROWS;
    a.n_tx = (a.geom.res_w + 63) / 64;
    a.n_ty = (a.geom.res_h + cap - 1) / cap;
    a.chunk_h = (a.geom.res_h + a.n_ty - 1) / a.n_ty;
    a.n_pairs = (a.tmpl_h + 1) / 2;
    a.kpp = (a.tmpl_w + 63 + 31) / 32;
    {
        static const bool all8 = std::getenv("MELF_TC_KPP8") && std::atoi(std::getenv("MELF_TC_KPP8"));   // A/B switch
        if (all8 || a.kpp > 8) a.kpp = 8;
    }
    a.n_items = (long long)n_frames * a.n_tx * a.n_ty;
    for (int j = 0; j < 64; ++j) {
        uint32_t w = 0;
        for (int b = 0; b < 4; ++b)
            if (4 * j + b < a.tmpl_w) w |= 1u << (8 * b);
        a.ones_words[j] = w;
    }
    return (a.n_items + 1) / 2;
}

size_t melf_match_tc_scratch_bytes(long long max_items, size_t *wsum_offset)
{
    const size_t slabs = (size_t)max_items * SLAB_BYTES;
    if (wsum_offset) *wsum_offset = slabs;
    return slabs + (size_t)max_items * WSUM_WORDS * sizeof(uint32_t);
}

// prep, then correlation, for the items [args.item0, args.item0 + args.max_items) of the batch
cudaError_t melf_launch_match_prep(const MelfMatchTcArgs &args, cudaStream_t stream)
{
    cudaError_t e = cudaFuncSetAttribute(melf_match_prep_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, PREP_SMEM_TOTAL);
    if (e != cudaSuccess) return e;
    long long n_local = args.n_items - args.item0;
    if (n_local > args.max_items) n_local = args.max_items;
    if (n_local <= 0) return cudaSuccess;
    melf_match_prep_kernel<<<(unsigned)n_local, PREP_THREADS, PREP_SMEM_TOTAL, stream>>>(args);
    return cudaGetLastError();
}

cudaError_t melf_launch_match_tc(const MelfMatchTcArgs &args, int num_sms, cudaStream_t stream)
{
    cudaError_t e = cudaFuncSetAttribute(melf_match_tc_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL2);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(melf_match_tc_kernel<true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL2);
    if (e != cudaSuccess) return e;
    long long n_local = args.n_items - args.item0;
    if (n_local > args.max_items) n_local = args.max_items;
    if (n_local <= 0) return cudaSuccess;
    const long long n_pass = (n_local + 1) / 2;
    const unsigned grid = (unsigned)(n_pass < num_sms ? n_pass : num_sms);
    if (args.result)
        melf_match_tc_kernel<true><<<grid, TC_THREADS, SMEM_TOTAL2, stream>>>(args);
    else
        melf_match_tc_kernel<false><<<grid, TC_THREADS, SMEM_TOTAL2, stream>>>(args);
    return cudaGetLastError();
}

// ---- host-only debug exports (tests/test_host_tables.py; need no GPU) -----------------------
extern "C" long long melf_host_match_tc_toeplitz(const uint8_t *tmpl, int tmpl_w, int tmpl_h,
                                                 uint8_t *out, long long cap)
{
    if (!melf_match_tc_supported(tmpl_w, tmpl_h)) return -1;
    const long long need = (long long)melf_match_tc_toeplitz_bytes(tmpl_h);
    if (out && cap >= need) melf_match_tc_build_toeplitz(tmpl, tmpl_w, tmpl_h, out);
    return need;
}

// plan[0..3] = n_tx, n_ty, chunk_h, n_pairs for a res_w x res_h result map
extern "C" int melf_host_match_tc_plan(int res_w, int res_h, int tmpl_w, int tmpl_h, int32_t *plan)
{
    if (!melf_match_tc_supported(tmpl_w, tmpl_h) || res_w < 1 || res_h < 1) return -1;
    MelfMatchTcArgs a{};
    a.geom.res_w = res_w; a.geom.res_h = res_h;
    a.tmpl_w = tmpl_w; a.tmpl_h = tmpl_h;
    melf_match_tc_plan(a, 1);
    plan[0] = a.n_tx; plan[1] = a.n_ty; plan[2] = a.chunk_h; plan[3] = a.n_pairs;
    return 0;
}

// =============================================================================================
// Small result maps: result positions as M, frames as N.
//
// With the Toeplitz formulation above one frame costs 60 x 8 MMAs whatever its result map: for sample set
// 2 (meter_rect 220 x 135, 33 x 17 = 561 positions) they run at N = 32, i.e. at the instruction's fixed
// cost (86 clk for the 4 KB Toeplitz tile) -- 6 % of the tensor peak.  When the whole map
// has few enough positions the correlation is one plain GEMM per group of 32 frames:
//     D[pos][frame] = sum_k Tpos[pos][k] * L_frame[k],   k = (r, c) over the crop,  Tpos[(y, x)][(r, c)] = T(r - y, c - x)
// M = positions (5 tiles of 128), N = 32 frames, K = crop_h x 224 = 30240: 945 K-steps x 5 tiles = 4725 MMAs per
// 32 frames instead of 480 per frame (3.3 x fewer), 75 % of the template operand non-zero.  Tpos (19 MB for
// set 2, built once per crop geometry) streams from L2, 40 KB per stage.
//
// One kernel, melf_small_fused_kernel, one CTA per group of 32 frames (the version with a stand-alone
// conversion kernel in front lost what the GEMM won: 0.35 + 0.26 ms against 0.53 ms, profiles/
// r02_small_launches.csv): ten worker warps convert the group's crops straight into the frame operand's
// shared-memory ring (lane = frame, 64 K bytes of all 32 frames per stage), in the shadow of the MMAs.
// Window sums: S(y, x) = C(x) - E(y, x) with C(x) = the box sum over ALL crop rows, which res_w spare rows of
// Tpos (ones over columns x .. x + tw - 1) deliver as extra accumulator lanes, and E(y, x) = the part of it
// above row y and from row y + th on -- 2 (res_h - 1) crop rows in all, summed by a small kernel in front
// (melf_small_edge_kernel).  Exact integer arithmetic like everything else here: bit-identical keys.
// =============================================================================================
namespace {

constexpr int SM_N = 32;                               // frames per group = MMA N
constexpr int SM_KS = 2;                               // K-steps of 32 per pipeline stage
constexpr int SM_MT_MAX = 5;
constexpr int SM_A_TILE = 2 * BSTRIP_BYTES;            // one K-step of one position tile: 4096
constexpr int SM_B_STEP = 2 * SM_N * 16;               // one K-step of the frame operand: 1024
constexpr int SM_B_STAGE = SM_KS * SM_B_STEP;          // 2048
constexpr int SF_THREADS = 512;
#ifndef MELF_SF_A_STAGES
#define MELF_SF_A_STAGES 3      // measured per 4096 set-2 frames: 3 + 32 stages 0.315 ms, 4 + 16 stages 0.330 ms, 3 + 48 stages 0.35 ms
#define MELF_SF_B_STAGES 32
#endif
constexpr int SF_A_STAGES = MELF_SF_A_STAGES;          // template ring (TMA), 40 KB per stage
constexpr int SF_B_STAGES = MELF_SF_B_STAGES;          // frame ring (written by the worker warps), 2 KB per stage
constexpr int SF_WORKER_WARP0 = 2, SF_WORKERS = 14;    // warp 0 TMA; 1 MMA issue; 2..15 convert, then 2..5 score
constexpr int SF_CBOX_MAX = 64;                        // res_w <= 64
#ifndef MELF_EDGE_THREADS
#define MELF_EDGE_THREADS 256   // (128 threads x 16 rows per round: 0.079 ms instead of 0.072 per 4096 frames)
#define MELF_EDGE_ROWS 32
#endif
constexpr int SM_EDGE_THREADS = MELF_EDGE_THREADS;
constexpr int SM_EDGE_ROWS = MELF_EDGE_ROWS;            // crop rows converted per round

struct CtrlSF {
    unsigned long long a_full[SF_A_STAGES];
    unsigned long long a_empty[SF_A_STAGES];
    unsigned long long b_full[SF_B_STAGES];
    unsigned long long b_empty[SF_B_STAGES];
    unsigned long long accum;
    unsigned long long keys[SM_N];
    uint32_t tmem_base;
    uint32_t pad;
};

// shared-memory pitches of the edge kernel, chosen so that its accesses do not pile up on a few banks
__host__ __device__ inline int sm_raw_pitch(int channels, int crop_w)
{
    int p = ((channels * crop_w + 15 + 15) & ~15) + 16;         // a row at any 16-byte phase
    if (((p >> 4) & 1) == 0) p += 16;                           // odd number of 16-byte chunks: 8 rows, 8 bank groups
    return p;
}
__host__ __device__ inline int sm_l_pitch(int cwp) { return cwp + 4; }       // odd number of words per row
__host__ __device__ inline int sm_edge_smem(int channels, int crop_w, int cwp, int res_w, int res_h)
{
    const int n_edge = 2 * (res_h - 1) > 0 ? 2 * (res_h - 1) : 1;
    return ((n_edge * sm_l_pitch(cwp) + 15) & ~15) + SM_EDGE_ROWS * sm_raw_pitch(channels, crop_w) + n_edge * (res_w | 1) * 4;
}

// 16 pixels of a crop row (starting at byte `pos` of the 16-byte-aligned staging row `rp`... see callers) ->
// 16 lightness bytes, pixels from `left` on set to zero
__device__ __forceinline__ uint4 sm_convert16(const uint32_t (&e)[17], uint32_t pos, int channels, int left)
{
    const uint32_t ws = (pos & 15u) >> 2, bs = (pos & 3u) * 8;
    uint32_t v[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        const uint32_t lo = ws == 0 ? e[j] : ws == 1 ? e[j + 1] : ws == 2 ? e[j + 2] : e[j + 3];
        const uint32_t hi = ws == 0 ? e[j + 1] : ws == 1 ? e[j + 2] : ws == 2 ? e[j + 3] : e[j + 4];
        v[j] = __funnelshift_r(lo, hi, bs);
    }
    const uint4 conv = channels == 1 ? make_uint4(v[0], v[1], v[2], v[3]) : lightness16(v);
    uint32_t w4[4] = {conv.x, conv.y, conv.z, conv.w};
    if (left < 16) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int keep = left - 4 * k;                              // bytes of word k that are pixels
            w4[k] = keep >= 4 ? w4[k] : keep <= 0 ? 0u : (w4[k] & ((1u << (8 * keep)) - 1u));
        }
    }
    return make_uint4(w4[0], w4[1], w4[2], w4[3]);
}

}  // namespace

// One CTA per frame: the 2 (res_h - 1) crop rows above the lowest window top and below the highest window bottom
// -> lightness -> horizontal box sums H(r, x) -> E(y, x) = sum_{r < y} H + sum_{r >= y + th} H, for every position.
__global__ void __launch_bounds__(SM_EDGE_THREADS)
melf_small_edge_kernel(const MelfSmallArgs A)
{
    extern __shared__ __align__(16) uint8_t sp_smem[];
    const int tid = threadIdx.x;
    const int f = blockIdx.x;
    const int cw = A.geom.crop_w, cwp = A.cwp;
    const int rw = A.geom.res_w, rh = A.geom.res_h;
    const int n_top = rh - 1, n_edge = 2 * n_top;                           // edge row e: crop row e (e < n_top) or th + e - n_top
    const int lp = sm_l_pitch(cwp);
    uint8_t *L = sp_smem;                                                   // [n_edge][lp]
    const int raw_pitch = sm_raw_pitch(A.channels, cw);
    uint8_t *raw = sp_smem + (((n_edge > 0 ? n_edge : 1) * lp + 15) & ~15);  // [SM_EDGE_ROWS][raw_pitch]
    const int hp = rw | 1;
    uint32_t *H = reinterpret_cast<uint32_t *>(raw + SM_EDGE_ROWS * raw_pitch);   // [n_edge][hp]
    const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
    const uint8_t *base = A.img + (size_t)f * A.frame_stride + (size_t)A.geom.crop_y0 * A.row_stride +
                          (size_t)A.channels * (size_t)A.geom.crop_x0;
    const int chunks = raw_pitch >> 4;
    const int units = cwp >> 4;
    auto crop_row = [&](int e) { return e < n_top ? e : A.tmpl_h + e - n_top; };
    for (int e0 = 0; e0 < n_edge; e0 += SM_EDGE_ROWS) {
        const int nr = min(SM_EDGE_ROWS, n_edge - e0);
        for (int i = tid; i < nr * chunks; i += SM_EDGE_THREADS) {
            const int r = i / chunks, c = i - r * chunks;
            const uintptr_t src = reinterpret_cast<uintptr_t>(base + (size_t)crop_row(e0 + r) * A.row_stride);
            const uintptr_t a16 = (src & ~(uintptr_t)15) + 16u * (uintptr_t)c;
            uint4 q = make_uint4(0, 0, 0, 0);
            if (a16 + 16 <= limit) q = __ldg(reinterpret_cast<const uint4 *>(a16));
            *reinterpret_cast<uint4 *>(raw + r * raw_pitch + 16 * c) = q;
        }
        __syncthreads();
        for (int i = tid; i < nr * units; i += SM_EDGE_THREADS) {
            const int u = i / nr, r = i - u * nr;                           // consecutive threads: consecutive rows
            const uintptr_t src = reinterpret_cast<uintptr_t>(base + (size_t)crop_row(e0 + r) * A.row_stride);
            const uint32_t pos = (uint32_t)(src & 15) + (uint32_t)(16 * A.channels * u);
            const uint8_t *rp = raw + r * raw_pitch + (pos & ~15u);
            uint32_t e[17];
#pragma unroll
            for (int cidx = 0; cidx < 4; ++cidx) {
                uint4 q = make_uint4(0, 0, 0, 0);
                if ((int)(pos >> 4) + cidx < chunks) q = *reinterpret_cast<const uint4 *>(rp + 16 * cidx);
                e[4 * cidx] = q.x; e[4 * cidx + 1] = q.y; e[4 * cidx + 2] = q.z; e[4 * cidx + 3] = q.w;
            }
            e[16] = 0;
            const uint4 out = sm_convert16(e, pos, A.channels, cw - 16 * u);
            uint32_t *lrow = reinterpret_cast<uint32_t *>(L + (size_t)(e0 + r) * lp + 16 * u);
            lrow[0] = out.x; lrow[1] = out.y; lrow[2] = out.z; lrow[3] = out.w;
        }
        __syncthreads();
    }
    // H(r, x): thread = (edge row, one of 8 runs of x): the run's first box sum by dp4a, then sliding
    {
        const int run = (rw + 7) >> 3;
        for (int t = tid; t < n_edge * 8; t += SM_EDGE_THREADS) {
            const int r = t >> 3, x0 = (t & 7) * run, x1 = min(rw, x0 + run);
            if (x0 >= x1) continue;
            const uint8_t *row = L + (size_t)r * lp;
            uint32_t s = 0;
            int j = x0;
            for (; (j & 3) && j < x0 + A.tmpl_w; ++j) s += row[j];
            const uint32_t *row32 = reinterpret_cast<const uint32_t *>(row);
            for (; j + 4 <= x0 + A.tmpl_w; j += 4) s = __dp4a(row32[j >> 2], 0x01010101u, s);
            for (; j < x0 + A.tmpl_w; ++j) s += row[j];
            for (int x = x0; x < x1; ++x) {
                H[r * hp + x] = s;
                s += (uint32_t)row[x + A.tmpl_w < cwp ? x + A.tmpl_w : cwp - 1] - (uint32_t)row[x];
            }
        }
    }
    __syncthreads();
    const int g = f / SM_N, slot = f - g * SM_N;
    uint32_t *E = A.wsum + (size_t)g * (A.n_mt * 128) * SM_N + slot;
    for (int p = tid; p < rw * rh; p += SM_EDGE_THREADS) {
        const int y = p / rw, x = p - y * rw;
        uint32_t s = 0;
        for (int e = 0; e < y; ++e) s += H[e * hp + x];                     // crop rows 0 .. y - 1
        for (int e = n_top + y; e < n_edge; ++e) s += H[e * hp + x];        // crop rows y + th .. crop_h - 1
        E[(size_t)p * SM_N] = s;
    }
}

__global__ void __launch_bounds__(SF_THREADS, 1)
melf_small_fused_kernel(const MelfSmallArgs A)
{
    uint8_t *smem = melf_tc_smem;
    const int a_bytes = A.n_mt * SM_KS * SM_A_TILE;
    uint8_t *b_ring = smem + SF_A_STAGES * a_bytes;
    uint32_t *cbox = reinterpret_cast<uint32_t *>(b_ring + SF_B_STAGES * SM_B_STAGE);      // [res_w][32 frames]
    CtrlSF *ctrl = reinterpret_cast<CtrlSF *>(reinterpret_cast<uint8_t *>(cbox) + SF_CBOX_MAX * SM_N * 4);
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int g = blockIdx.x;
    if (tid == 0) {
        for (int s = 0; s < SF_A_STAGES; ++s) { mbar_init(smem_u32(&ctrl->a_full[s]), 1); mbar_init(smem_u32(&ctrl->a_empty[s]), 1); }
        for (int s = 0; s < SF_B_STAGES; ++s) { mbar_init(smem_u32(&ctrl->b_full[s]), 1); mbar_init(smem_u32(&ctrl->b_empty[s]), 1); }
        mbar_init(smem_u32(&ctrl->accum), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < SM_N) ctrl->keys[tid] = 0ull;
    __syncwarp();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;"
                     :: "r"(smem_u32(&ctrl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctrl->tmem_base;
    const uint32_t a_ring_s = smem_u32(smem), b_ring_s = smem_u32(b_ring);
    const int n_valid = min(SM_N, A.n_frames - g * SM_N);

    if (warp == 0) {
        // ================= template tiles: TMA, 3-stage ring =================
        if (elect_one()) {
            const u64 pol_keep = policy_evict_last();
            for (int st = 0; st < A.n_kstages; ++st) {
                const uint32_t slot = (uint32_t)st % SF_A_STAGES, ph = ((uint32_t)st / SF_A_STAGES) & 1;
                mbar_wait(smem_u32(&ctrl->a_empty[slot]), ph ^ 1);
                mbar_expect_tx(smem_u32(&ctrl->a_full[slot]), (uint32_t)a_bytes);
                bulk_load_keep(a_ring_s + slot * a_bytes, A.tpos + (size_t)st * a_bytes, (uint32_t)a_bytes,
                               smem_u32(&ctrl->a_full[slot]), pol_keep);
            }
        }
    } else if (warp == 1) {
        // ================= MMA issue =================
        const bool lead = elect_one();
        const uint32_t idesc = make_idesc(SM_N);
        for (int st = 0; st < A.n_kstages; ++st) {
            const uint32_t sa = (uint32_t)st % SF_A_STAGES, pa = ((uint32_t)st / SF_A_STAGES) & 1;
            const uint32_t sb = (uint32_t)st % SF_B_STAGES, pb = ((uint32_t)st / SF_B_STAGES) & 1;
            mbar_wait(smem_u32(&ctrl->a_full[sa]), pa);
            mbar_wait(smem_u32(&ctrl->b_full[sb]), pb);
            tc_fence_after();
            if (lead) {
                const uint32_t abase = a_ring_s + sa * a_bytes, bbase = b_ring_s + sb * SM_B_STAGE;
                for (int mt = 0; mt < A.n_mt; ++mt) {
                    if (st < A.mt_k0[mt] || st >= A.mt_k1[mt]) continue;     // all-zero template bytes for this tile
#pragma unroll
                    for (int ks = 0; ks < SM_KS; ++ks)
                        umma_i8(tmem + (uint32_t)(mt * SM_N), umma_desc(abase + (mt * SM_KS + ks) * SM_A_TILE, BSTRIP_BYTES),
                                umma_desc(bbase + ks * SM_B_STEP, SM_N * 16), idesc, (st > A.mt_k0[mt] || ks > 0) ? 1u : 0u);
                }
                umma_commit(smem_u32(&ctrl->a_empty[sa]));
                umma_commit(smem_u32(&ctrl->b_empty[sb]));
            }
            __syncwarp();
        }
        if (lead) umma_commit(smem_u32(&ctrl->accum));
    } else {
        // ================= frame operand: lane = frame, one stage (64 K bytes of all 32 frames) per turn =================
        const int ww = warp - SF_WORKER_WARP0;
        const int f = g * SM_N + lane;
        const bool have = f < A.n_frames;
        const uintptr_t limit = reinterpret_cast<uintptr_t>(A.img_limit);
        const uint8_t *base = A.img + (size_t)(have ? f : 0) * A.frame_stride + (size_t)A.geom.crop_y0 * A.row_stride +
                              (size_t)A.channels * (size_t)A.geom.crop_x0;
        const int units = A.cwp >> 4;
        const int nchunk = A.channels == 3 ? 4 : 2;
        for (int st = ww; st < A.n_kstages; st += SF_WORKERS) {
            const uint32_t sb = (uint32_t)st % SF_B_STAGES, pb = ((uint32_t)st / SF_B_STAGES) & 1;
            mbar_wait(smem_u32(&ctrl->b_empty[sb]), pb ^ 1);
            uint8_t *bst = b_ring + sb * SM_B_STAGE + lane * 16;
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                uint32_t e[2][17];
                uint32_t pos2[2];
                int left2[2];
                bool live2[2];
                // the loads of both units first (two 64-byte windows in flight per lane), then the conversions
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int k16 = 4 * st + 2 * half + q;
                    const int row = k16 / units, u = k16 - row * units;
                    live2[q] = have && row < A.geom.crop_h && 16 * u < A.geom.crop_w;
                    left2[q] = A.geom.crop_w - 16 * u;
                    const uintptr_t src = reinterpret_cast<uintptr_t>(base + (size_t)row * A.row_stride) + (uintptr_t)(16 * A.channels * u);
                    pos2[q] = (uint32_t)(src & 15);
                    const uint4 *al = reinterpret_cast<const uint4 *>(src & ~(uintptr_t)15);
#pragma unroll
                    for (int cidx = 0; cidx < 4; ++cidx) {
                        uint4 v = make_uint4(0, 0, 0, 0);
                        if (live2[q] && cidx < nchunk && reinterpret_cast<uintptr_t>(al + cidx) + 16 <= limit) v = __ldg(al + cidx);
                        e[q][4 * cidx] = v.x; e[q][4 * cidx + 1] = v.y; e[q][4 * cidx + 2] = v.z; e[q][4 * cidx + 3] = v.w;
                    }
                    e[q][16] = 0;
                }
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    uint4 out = make_uint4(0, 0, 0, 0);
                    if (live2[q]) out = sm_convert16(e[q], pos2[q], A.channels, left2[q]);
                    // unit j of the stage = K-step j >> 1, strip j & 1: [K-step][strip][frame][16 B]
                    *reinterpret_cast<uint4 *>(bst + (2 * half + q) * (SM_N * 16)) = out;
                }
            }
            fence_async_smem();                    // generic-proxy writes before the MMA's async-proxy reads
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&ctrl->b_full[sb]));
        }
    }
    if (warp >= 2 && warp < 6) {
        // ================= scores and arg-max per frame (four of the worker warps, one per TMEM lane quadrant) =================
        const int qd = warp & 3;
        const int et = tid - 64;                                           // 0 .. 127
        const int rw = A.geom.res_w;
        mbar_wait(smem_u32(&ctrl->accum), 0);
        tc_fence_after();
        const uint32_t n_t32 = (uint32_t)A.n_t, sum_t32 = (uint32_t)A.sum_t;
        // the box sums over all rows: accumulator lanes n_pos .. n_pos + res_w - 1, to shared memory
        for (int mt = A.n_pos >> 7; mt < A.n_mt; ++mt) {
            const int pos = mt * 128 + 32 * qd + lane;
            uint32_t d[33];
            tmem_ld32(tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(mt * SM_N), d);
            tmem_ld_wait();
            if (pos >= A.n_pos && pos < A.n_pos + rw) {
#pragma unroll
                for (int k = 0; k < SM_N; ++k) cbox[(pos - A.n_pos) * SM_N + k] = d[k];
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // best key of a frame over the 32 positions of this warp, then one shared-memory atomic per warp
        auto q4_keys = [&](int q4, u64 (&k4)[4]) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                u64 best = k4[k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
                    best = other > best ? other : best;
                }
                if (lane == 0 && 4 * q4 + k < n_valid && best) atomicMax(&ctrl->keys[4 * q4 + k], best);
            }
        };
        const int n_score_mt = (A.n_pos + 127) >> 7;
        for (int mt = 0; mt < n_score_mt; ++mt) {
            const int pos = mt * 128 + 32 * qd + lane;
            u64 keys4[4];
            uint32_t d[33];
            tmem_ld32(tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(mt * SM_N), d);
            tmem_ld_wait();
            if (pos < A.n_pos) {
                const int x = pos % rw;
                const uint4 *E = reinterpret_cast<const uint4 *>(A.wsum + ((size_t)g * (A.n_mt * 128) + pos) * SM_N);
                const uint4 *C = reinterpret_cast<const uint4 *>(cbox + x * SM_N);
#pragma unroll
                for (int q4 = 0; q4 < SM_N / 4; ++q4) {
                    const uint4 ew = __ldcg(E + q4), cwv = C[q4];
                    const uint32_t s4[4] = {cwv.x - ew.x, cwv.y - ew.y, cwv.z - ew.z, cwv.w - ew.w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const long long num = (long long)((u64)n_t32 * (u64)d[4 * q4 + k]) - (long long)((u64)sum_t32 * (u64)s4[k]);
                        keys4[k] = melf_score_key(score_of(num, A.n_t), (uint32_t)pos);
                    }
                    q4_keys(q4, keys4);
                }
            } else {
#pragma unroll
                for (int q4 = 0; q4 < SM_N / 4; ++q4) {
                    u64 none[4] = {0ull, 0ull, 0ull, 0ull};
                    q4_keys(q4, none);
                }
            }
        }
        tc_fence_before();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et < n_valid) A.match_key[(size_t)g * SM_N + et] = ctrl->keys[et];
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tmem) : "memory");
}

// ---- host side of the small-map path ---------------------------------------------------------------------
bool melf_match_small_supported(const MelfGeom &g, int tmpl_w, int tmpl_h, int channels)
{
    const long long n_pos = (long long)g.res_w * g.res_h;
    const int cwp = (g.crop_w + 31) & ~31;
    const long long k_total = (long long)g.crop_h * cwp;
    return n_pos + g.res_w <= SM_MT_MAX * 128 && g.res_w <= SF_CBOX_MAX && tmpl_w <= g.crop_w && tmpl_h <= g.crop_h &&
           k_total <= 64 * 1024 && g.res_h <= 64 &&
           sm_edge_smem(channels, g.crop_w, cwp, g.res_w, g.res_h) <= 200 * 1024;
}

void melf_match_small_plan(MelfSmallArgs &a)
{
    a.n_pos = a.geom.res_w * a.geom.res_h;
    const int n_rows = a.n_pos + a.geom.res_w;                       // + the box-sum rows
    a.n_mt = (n_rows + 127) / 128;
    a.cwp = (a.geom.crop_w + 31) & ~31;
    a.n_kstages = (a.geom.crop_h * a.cwp + 32 * SM_KS - 1) / (32 * SM_KS);
    for (int mt = 0; mt < SM_MT_MAX; ++mt) { a.mt_k0[mt] = 0; a.mt_k1[mt] = 0; }
    for (int mt = 0; mt < a.n_mt; ++mt) {
        const int p0 = mt * 128, p1 = (p0 + 128 < n_rows ? p0 + 128 : n_rows) - 1;
        if (p1 >= a.n_pos) { a.mt_k0[mt] = 0; a.mt_k1[mt] = a.n_kstages; continue; }   // a box-sum row: every crop row
        const int y0 = p0 / a.geom.res_w, y1 = p1 / a.geom.res_w;
        const long long k_lo = (long long)y0 * a.cwp, k_hi = (long long)(y1 + a.tmpl_h) * a.cwp;
        a.mt_k0[mt] = (int)(k_lo / (32 * SM_KS));
        a.mt_k1[mt] = (int)((k_hi + 32 * SM_KS - 1) / (32 * SM_KS));
        if (a.mt_k1[mt] > a.n_kstages) a.mt_k1[mt] = a.n_kstages;
    }
}

size_t melf_match_small_tpos_bytes(const MelfSmallArgs &a) { return (size_t)a.n_kstages * a.n_mt * SM_KS * SM_A_TILE; }
size_t melf_match_small_wsum_bytes(const MelfSmallArgs &a)
{
    return (size_t)((a.n_frames + SM_N - 1) / SM_N) * (a.n_mt * 128) * SM_N * 4;
}

// Tpos: [K stage][position tile][K-step][strip][128 rows][16 B]; row = position (y, x): byte = T(r - y, c - x) for
// K index (r, c); row n_pos + x: 1 for every crop row r and x <= c < x + tw (the box sum C(x))
void melf_match_small_build_tpos(const MelfSmallArgs &a, const uint8_t *tmpl, uint8_t *out)
{
    std::memset(out, 0, melf_match_small_tpos_bytes(a));
    auto at = [&](int row, long long k) -> uint8_t & {
        const long long kstep = k >> 5;
        const int st = (int)(kstep / SM_KS), ks = (int)(kstep % SM_KS), strip = (int)((k >> 4) & 1), b = (int)(k & 15);
        const int mt = row >> 7, m = row & 127;
        return out[((((size_t)st * a.n_mt + mt) * SM_KS + ks) * 2 + strip) * BSTRIP_BYTES + (size_t)m * 16 + b];
    };
    for (int pos = 0; pos < a.n_pos; ++pos) {
        const int y = pos / a.geom.res_w, x = pos - y * a.geom.res_w;
        for (int i = 0; i < a.tmpl_h; ++i)
            for (int j = 0; j < a.tmpl_w; ++j) at(pos, (long long)(y + i) * a.cwp + (x + j)) = tmpl[(size_t)i * a.tmpl_w + j];
    }
    for (int x = 0; x < a.geom.res_w; ++x)
        for (int r = 0; r < a.geom.crop_h; ++r)
            for (int j = 0; j < a.tmpl_w; ++j) at(a.n_pos + x, (long long)r * a.cwp + (x + j)) = 1;
}

cudaError_t melf_launch_match_small(const MelfSmallArgs &a, cudaStream_t stream)
{
    const int edge_smem = sm_edge_smem(a.channels, a.geom.crop_w, a.cwp, a.geom.res_w, a.geom.res_h);
    cudaError_t e = cudaFuncSetAttribute(melf_small_edge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, edge_smem);
    if (e != cudaSuccess) return e;
    const int fused_smem = SF_A_STAGES * (a.n_mt * SM_KS * SM_A_TILE) + SF_B_STAGES * SM_B_STAGE + SF_CBOX_MAX * SM_N * 4 +
                           (int)sizeof(CtrlSF);
    e = cudaFuncSetAttribute(melf_small_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, fused_smem);
    if (e != cudaSuccess) return e;
    if (a.n_frames <= 0) return cudaSuccess;
    melf_small_edge_kernel<<<a.n_frames, SM_EDGE_THREADS, edge_smem, stream>>>(a);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    melf_small_fused_kernel<<<(a.n_frames + SM_N - 1) / SM_N, SF_THREADS, fused_smem, stream>>>(a);
    return cudaGetLastError();
}
