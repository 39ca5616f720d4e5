// Dial location on the 5th-generation tensor cores: TM_CCOEFF as an exact u8 x u8 -> s32
// implicit GEMM (tcgen05.mma.kind::i8, accumulators in TMEM), fused with the lightness
// conversion in front of it and the window sums, score and arg-max behind it.
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
// One CTA (512 threads) per SM, persistent over "passes" of two work items (a work item
// = one 64-column x <=137-row block of one frame's result map; for the sample geometry,
// 63 x 132, that is one whole frame).  Per pass:
//   1. all warps: BGR crop -> lightness (bit-exact float32 order of cv2, no XU pipe ops)
//      -> two 64 KB slabs in shared memory;
//   2. one thread streams the Toeplitz pairs (bulk async copies + mbarriers), one thread
//      issues 60 x (2 items x 2 M-tiles x 8 K-steps) tcgen05.mma of 128 x 128 x 32;
//   3. all warps: horizontal then vertical window sums of the slabs (the sum_window L
//      term), into the now idle operand buffers;
//   4. all warps: TMEM -> registers, exact score (N*sumTL - sumT*sumL)/N -> float32,
//      64-bit (score, ~raster index) key, warp/block max, one atomicMax per item.
// Everything is integer-exact, so results are bit-identical to melf_match.cu and to the
// oracle.
#include "melf_common.cuh"
#include "melf_args.h"

namespace {

constexpr int TC_THREADS = 512;
constexpr int SLAB_ROWS = 256;                       // rows of a lightness slab
constexpr int KS = 16;                               // 16-byte K strips: K = 256
constexpr int STRIP_BYTES = SLAB_ROWS * 16;          // 4096
constexpr int SLAB_BYTES = KS * STRIP_BYTES;         // 65536
constexpr int BN = 128;                              // MMA N: two template rows x 64 columns
constexpr int BSTRIP_BYTES = BN * 16;                // 2048
constexpr int BSTAGE_BYTES = KS * BSTRIP_BYTES;      // 32768
constexpr int NSTAGE = 2;
constexpr int SROWS = 137;                           // max result rows of a work item (odd)
constexpr int SMEM_SLABS = 0;
constexpr int SMEM_B = 2 * SLAB_BYTES;               // 131072
constexpr int SMEM_CTRL = SMEM_B + NSTAGE * BSTAGE_BYTES;   // 196608
constexpr int SMEM_TOTAL = SMEM_CTRL + 512;

// tcgen05 instruction descriptor: D = s32, A = B = u8, both K-major, M = 128, N = 128
constexpr uint32_t IDESC = (2u << 4) | (0u << 7) | (0u << 10) | ((BN >> 3) << 17) | ((128u >> 4) << 24);

struct Ctrl {
    unsigned long long full[NSTAGE];
    unsigned long long empty[NSTAGE];
    unsigned long long accum;
    uint32_t tmem_base;
    uint32_t pad;
    u64 red[16];
};

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

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, u64 adesc, u64 bdesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}" :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(IDESC), "r"(accumulate) : "memory");
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
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// geometry of one work item
struct Item {
    int frame, x0, y0, w_c, h_c, off1, n_mt;
    bool valid;
};

__device__ __forceinline__ Item get_item(const MelfMatchTcArgs &A, long long idx)
{
    Item it;
    it.valid = idx < A.n_items;
    if (!it.valid) { it.frame = 0; it.x0 = it.y0 = 0; it.w_c = it.h_c = 0; it.off1 = 0; it.n_mt = 0; return it; }
    const int per_frame = A.n_tx * A.n_ty;
    it.frame = (int)(idx / per_frame);
    const int rem = (int)(idx - (long long)it.frame * per_frame);
    const int ty = rem / A.n_tx, tx = rem - ty * A.n_tx;
    it.x0 = tx * 64;
    it.y0 = ty * A.chunk_h;
    it.w_c = min(64, A.geom.res_w - it.x0);
    it.h_c = min(A.chunk_h, A.geom.res_h - it.y0);
    it.n_mt = it.h_c > 127 ? 2 : 1;
    it.off1 = it.h_c - 127;                    // M-tile 1 covers rows [h_c-127, h_c)
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

template <bool WRITE_MAP>
__global__ void __launch_bounds__(TC_THREADS, 1)
melf_match_tc_kernel(const MelfMatchTcArgs A)
{
    uint8_t *smem = melf_tc_smem;
    Ctrl *ctrl = reinterpret_cast<Ctrl *>(smem + SMEM_CTRL);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t slab_addr = smem_u32(smem + SMEM_SLABS), b_addr = smem_u32(smem + SMEM_B);

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(smem_u32(&ctrl->full[s]), 1);
            mbar_init(smem_u32(&ctrl->empty[s]), 1);
        }
        mbar_init(smem_u32(&ctrl->accum), 1);
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

    uint32_t stage_count = 0;       // Toeplitz stages consumed so far (producer and issuer keep their own copy)
    uint32_t pass_count = 0;
    const long long n_pass = (A.n_items + 1) / 2;

    for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x, ++pass_count) {
        Item items[2];
        items[0] = get_item(A, 2 * pass);
        items[1] = get_item(A, 2 * pass + 1);

        // ---- 1. lightness slabs -----------------------------------------------------------
        // warp-unit = 8 rows x 4 strips; lane = strip * 8 + row, so that every quarter warp
        // stores 128 contiguous bytes (conflict-free 128-bit shared stores)
        for (int u = warp; u < 2 * 32 * 4; u += TC_THREADS / 32) {
            const int slot = u >> 7, rg = (u >> 2) & 31, sg = u & 3;
            const int r = rg * 8 + (lane & 7), s = sg * 4 + (lane >> 3);
            const uint4 v = stage_unit(A, slot ? items[1] : items[0], r, s);
            *reinterpret_cast<uint4 *>(smem + SMEM_SLABS + slot * SLAB_BYTES + s * STRIP_BYTES + r * 16) = v;
        }
        fence_async_smem();          // generic-proxy writes -> visible to the tensor core / bulk copies
        tc_fence_before();
        __syncthreads();
        tc_fence_after();

        // ---- 2. the correlation: Toeplitz stream + MMA -----------------------------------
        if (tid == 0) {
            uint32_t g = stage_count;
            for (int p = 0; p < A.n_pairs; ++p, ++g) {
                const uint32_t st = g & 1, ph = (g >> 1) & 1;
                mbar_wait(smem_u32(&ctrl->empty[st]), ph ^ 1);
                mbar_expect_tx(smem_u32(&ctrl->full[st]), BSTAGE_BYTES);
                bulk_load(b_addr + st * BSTAGE_BYTES, A.toeplitz + (size_t)p * BSTAGE_BYTES, BSTAGE_BYTES,
                          smem_u32(&ctrl->full[st]));
            }
        } else if (tid == 32) {
            uint32_t g = stage_count;
            for (int p = 0; p < A.n_pairs; ++p, ++g) {
                const uint32_t st = g & 1, ph = (g >> 1) & 1;
                mbar_wait(smem_u32(&ctrl->full[st]), ph);
                tc_fence_after();
                const uint32_t bbase = b_addr + st * BSTAGE_BYTES;
#pragma unroll
                for (int slot = 0; slot < 2; ++slot) {
                    if (!items[slot].valid) continue;
                    for (int mt = 0; mt < items[slot].n_mt; ++mt) {
                        const int row0 = (mt ? items[slot].off1 : 0) + 2 * p;
                        const uint32_t abase = slab_addr + slot * SLAB_BYTES + row0 * 16;
                        const uint32_t d = tmem + (uint32_t)((slot * 2 + mt) * BN);
#pragma unroll
                        for (int ks = 0; ks < KS / 2; ++ks)
                            umma_i8(d, umma_desc(abase + ks * 2 * STRIP_BYTES, STRIP_BYTES),
                                    umma_desc(bbase + ks * 2 * BSTRIP_BYTES, BSTRIP_BYTES),
                                    (p > 0 || ks > 0) ? 1u : 0u);
                    }
                }
                umma_commit(smem_u32(&ctrl->empty[st]));
            }
            umma_commit(smem_u32(&ctrl->accum));
            mbar_wait(smem_u32(&ctrl->accum), pass_count & 1);
        }
        stage_count += (uint32_t)A.n_pairs;
        tc_fence_before();
        __syncthreads();
        tc_fence_after();

        // ---- 3a. horizontal window sums H(row, x) = sum_{j<tw} L(row, x+j), x < 64 ----------
        // thread = (slot, row); stored as u16 pairs, word (row, xp) at row*32 + (xp ^ (row & 31))
        {
            const int slot = tid >> 8, row = tid & 255;
            const uint8_t *sl = smem + SMEM_SLABS + slot * SLAB_BYTES + row * 16;
            uint32_t *Hb = reinterpret_cast<uint32_t *>(smem + SMEM_B + slot * 32768) + row * 32;
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
        __syncthreads();

        // ---- 3b. vertical: S(y, x) = sum_{i<th} H(y+i, x); thread = (slot, segment, x pair) --
        // S goes into the (now dead) slab of the slot, x-major: S[x * SROWS + y]
        {
            const int slot = tid >> 8, seg = (tid >> 5) & 7, xp = lane;
            const Item it = slot ? items[1] : items[0];
            const uint32_t *Hb = reinterpret_cast<const uint32_t *>(smem + SMEM_B + slot * 32768);
            uint32_t *S = reinterpret_cast<uint32_t *>(smem + SMEM_SLABS + slot * SLAB_BYTES);
            const int rs = (it.h_c + 7) >> 3;
            const int ya = seg * rs, yb = min(it.h_c, ya + rs);
            if (ya < yb) {
                uint32_t s0 = 0, s1 = 0;
                for (int i = 0; i < A.tmpl_h; ++i) {
                    const int row = ya + i;
                    const uint32_t v = Hb[row * 32 + (xp ^ (row & 31))];
                    s0 += v & 0xffff; s1 += v >> 16;
                }
                for (int y = ya; y < yb; ++y) {
                    S[(2 * xp) * SROWS + y] = s0;
                    S[(2 * xp + 1) * SROWS + y] = s1;
                    const int rin = y + A.tmpl_h;
                    const uint32_t vin = Hb[rin * 32 + (xp ^ (rin & 31))];      // row <= 255 always
                    const uint32_t vout = Hb[y * 32 + (xp ^ (y & 31))];
                    s0 += (vin & 0xffff) - (vout & 0xffff);
                    s1 += (vin >> 16) - (vout >> 16);
                }
            }
        }
        __syncthreads();

        // ---- 4. drain: scores + arg-max ----------------------------------------------------
        const int qd = warp & 3, acc = warp >> 2, slot = acc >> 1, mt = acc & 1;
        const Item it = slot ? items[1] : items[0];
        const bool active = it.valid && mt < it.n_mt;
        const uint32_t taddr = tmem + ((uint32_t)(32 * qd) << 16) + (uint32_t)(acc * BN);
        uint32_t *xchg = reinterpret_cast<uint32_t *>(smem + SMEM_B);      // [16 warps][64]
        if (active) {
            // lane 0's D_1 row (tile row 32*qd) is the "row below" of the previous quadrant's lane 31
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t v[16];
                tmem_ld16(taddr + 64 + 16 * c, v);
                tmem_ld_wait();
                if (lane == 0)
#pragma unroll
                    for (int i = 0; i < 16; ++i) xchg[warp * 64 + 16 * c + i] = v[i];
            }
        }
        __syncthreads();
        u64 best = 0;
        if (active) {
            const int trow = 32 * qd + lane;
            const int yl = (mt ? it.off1 : 0) + trow;                  // row inside the work item
            const bool row_ok = trow < 127 && yl < it.h_c && (mt == 0 || yl >= 127);
            const uint32_t *S = reinterpret_cast<const uint32_t *>(smem + SMEM_SLABS + slot * SLAB_BYTES);
            const uint32_t row_index = (uint32_t)(it.y0 + yl) * (uint32_t)A.geom.res_w + (uint32_t)it.x0;
            float *map = WRITE_MAP ? A.result + (size_t)it.frame * A.geom.res_w * A.geom.res_h : nullptr;
            long long thr = -0x7fffffffffffffffll - 1;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t d0[16], d1[16];
                tmem_ld16(taddr + 16 * c, d0);
                tmem_ld16(taddr + 64 + 16 * c, d1);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int x = 16 * c + i;
                    uint32_t below = __shfl_down_sync(0xffffffffu, d1[i], 1);
                    if (lane == 31) below = qd < 3 ? xchg[(warp + 1) * 64 + x] : 0u;
                    if (row_ok && x < it.w_c) {
                        const long long sum_tl = (long long)(int)(d0[i] + below);
                        const long long sum_win = (long long)S[x * SROWS + yl];
                        const long long num = A.n_t * sum_tl - A.sum_t * sum_win;
                        if (WRITE_MAP || num >= thr) {
                            const float v = __double2float_rn(__ddiv_rn((double)num, (double)A.n_t));
                            if (WRITE_MAP) map[row_index + (uint32_t)x] = v;
                            const u64 key = melf_score_key(v, row_index + (uint32_t)x);
                            if (key > best) {
                                best = key;
                                // a later value can only tie or beat this one if it is within two
                                // float32 ulps of it; (|num| + n_t) >> 21 is 4x that, in num units
                                const long long mag = (num < 0 ? -num : num) + A.n_t;
                                thr = num - (mag >> 21) - 1;
                            }
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if (lane == 0) ctrl->red[warp] = best;
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        if (tid < 2) {
            const Item mine = tid ? items[1] : items[0];
            u64 b = 0;
            for (int w8 = 0; w8 < 8; ++w8) {
                const u64 k = ctrl->red[tid * 8 + w8];
                b = k > b ? k : b;
            }
            if (mine.valid && b) atomicMax(A.match_key + mine.frame, b);
        }
        // the next pass overwrites the slabs (S tables) and red[]: every reader is past the
        // barrier above except the two atomics, which only read red[]
        __syncthreads();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
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

// Toeplitz operand image: [pair p][strip s][n 0..127][16 bytes]; n < 64: template row 2p,
// n >= 64: template row 2p+1 (zero beyond the template); byte b of strip s is K index
// k = 16 s + b and holds T(row, k - (n mod 64)).
void melf_match_tc_build_toeplitz(const uint8_t *tmpl, int tmpl_w, int tmpl_h, uint8_t *out)
{
    const int n_pairs = (tmpl_h + 1) / 2;
    for (int p = 0; p < n_pairs; ++p)
        for (int s = 0; s < KS; ++s)
            for (int n = 0; n < BN; ++n)
                for (int b = 0; b < 16; ++b) {
                    const int row = 2 * p + (n >> 6), j = 16 * s + b - (n & 63);
                    uint8_t v = 0;
                    if (row < tmpl_h && j >= 0 && j < tmpl_w) v = tmpl[(size_t)row * tmpl_w + j];
                    out[(((size_t)p * KS + s) * BN + n) * 16 + b] = v;
                }
}

// fills the tiling fields of args; returns the number of passes
long long melf_match_tc_plan(MelfMatchTcArgs &a, int n_frames)
{
    const int max_h = SLAB_ROWS - a.tmpl_h;            // rows r + 2p <= h_c + th - 1 <= 255
    const int cap = max_h < SROWS ? max_h : SROWS;
    a.n_tx = (a.geom.res_w + 63) / 64;
    a.n_ty = (a.geom.res_h + cap - 1) / cap;
    a.chunk_h = (a.geom.res_h + a.n_ty - 1) / a.n_ty;
    a.n_pairs = (a.tmpl_h + 1) / 2;
    a.n_items = (long long)n_frames * a.n_tx * a.n_ty;
    for (int j = 0; j < 64; ++j) {
        uint32_t w = 0;
        for (int b = 0; b < 4; ++b)
            if (4 * j + b < a.tmpl_w) w |= 1u << (8 * b);
        a.ones_words[j] = w;
    }
    return (a.n_items + 1) / 2;
}

cudaError_t melf_launch_match_tc(const MelfMatchTcArgs &args, long long n_pass, int num_sms,
                                 cudaStream_t stream)
{
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(melf_match_tc_kernel<false>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(melf_match_tc_kernel<true>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    const unsigned grid = (unsigned)(n_pass < num_sms ? n_pass : num_sms);
    if (args.result)
        melf_match_tc_kernel<true><<<grid, TC_THREADS, SMEM_TOTAL, stream>>>(args);
    else
        melf_match_tc_kernel<false><<<grid, TC_THREADS, SMEM_TOTAL, stream>>>(args);
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
