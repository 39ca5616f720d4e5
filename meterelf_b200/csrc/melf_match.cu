// Dial location: TM_CCOEFF of the lightness plane against the dials template + argmax.
//
// Replaces ImageFile._find_dials (_image.py:57-66) = cv2.split(hls)[1] +
// match_template (_utils.py:91-97: cv2.matchTemplate(TM_CCOEFF) + cv2.minMaxLoc) for a
// whole batch of frames.
//
//   R(x,y) = sum_ij (T(i,j) - mean T) * L(y+i, x+j)
//          = (N_T * sum_ij T*L  -  sum T * sum_window L) / N_T
// evaluated EXACTLY: the u8 x u8 correlation in int32 (<= 1.46e9 for 188x119), the
// window sum in int32, the numerator in int64, one float64 division, one rounding to
// float32 (OpenCV's result dtype).  OpenCV itself goes through a float32 DFT and is up
// to ~2e2 away from this value at magnitudes of 2-4e7; the arg-max agrees on all sample
// frames.
//
// v1 kernel: direct correlation on the integer dot-product pipe (dp4a), staged through
// shared memory.  One CTA owns a tile of 64 x (NYT*8) result positions of one frame:
//   1. the lightness tile ((NYT*8 + th-1) rows x (64 + tw-1) columns) is computed from
//      the BGR crop on the fly (or copied from a 1-channel image) into shared memory as
//      packed bytes; the template sits next to it, pre-packed 4 columns per word and
//      zero padded;
//   2. each thread accumulates a register tile of 4 (x) x 8 (y) results: per step it
//      reads 3 lightness words and one template word pair, builds the 4 byte-shifted
//      windows with funnel shifts and issues 64 dp4a;
//   3. window sums come from a small separable pass (row sums in shared memory);
//   4. per-thread / warp-shuffle / block reduction of the 64-bit score key, one
//      atomicMax per CTA on the frame's slot.
#include "melf_common.cuh"
#include "melf_args.h"

namespace {

constexpr int RY = 8;          // result rows per thread
constexpr int XT = 16;         // x-threads per tile: 16 * 4 = 64 result columns

__device__ __forceinline__ int dp4a_u(uint32_t a, uint32_t b, int c)
{
    return (int)__dp4a(a, b, (uint32_t)c);
}

}  // namespace


// dynamic shared memory layout: lightness tile | template | row window sums
extern __shared__ __align__(16) uint8_t melf_match_smem[];

__global__ void __launch_bounds__(XT * 32)
melf_match_kernel(const MelfMatchArgs A)
{
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * XT + tx, nthreads = XT * A.nyt;
    const int f = blockIdx.z;
    const int tile_x0 = blockIdx.x * (XT * 4);
    const int tile_y0 = blockIdx.y * (A.nyt * RY);

    uint32_t *Ls = reinterpret_cast<uint32_t *>(melf_match_smem);
    uint32_t *Ts = Ls + (size_t)A.tile_rows * A.pitch_w;
    uint16_t *Hs = reinterpret_cast<uint16_t *>(Ts + (size_t)A.tw4p * A.steps);

    // ---- 1. stage the lightness tile (zero outside the crop) and the template ---------
    {
        const uint8_t *frame = A.img + (size_t)f * A.frame_stride;
        const int nwords = A.tile_rows * A.pitch_w;
        for (int i = tid; i < nwords; i += nthreads) {
            const int row = i / A.pitch_w, wc = i - row * A.pitch_w;
            const int cy = tile_y0 + row;
            uint32_t word = 0;
            if (cy < A.geom.crop_h) {
                const uint8_t *p = frame + (size_t)(A.geom.crop_y0 + cy) * A.row_stride;
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int cx = tile_x0 + wc * 4 + b;
                    if (cx < A.geom.crop_w) {
                        uint32_t v;
                        if (A.channels == 3) {
                            const uint8_t *q = p + 3 * (size_t)(A.geom.crop_x0 + cx);
                            v = (uint32_t)melf_lightness(q[0], q[1], q[2]);
                        } else {
                            v = p[A.geom.crop_x0 + cx];
                        }
                        word |= v << (8 * b);
                    }
                }
            }
            Ls[i] = word;
        }
        const int tmpl_n = A.tw4p * A.steps;
        for (int i = tid; i < tmpl_n; i += nthreads) Ts[i] = A.tmpl_words[i];
    }
    __syncthreads();

    // ---- 2. row window sums: Hs[row][x] = sum_j L[row][tile_x + x + j], x < 64 --------
    {
        const int ntasks = A.tile_rows * XT;
        for (int t = tid; t < ntasks; t += nthreads) {
            const int row = t / XT, g = t - row * XT;
            const uint32_t *Lp = Ls + (size_t)row * A.pitch_w + g;
            int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
            uint32_t w0 = Lp[0];
            for (int j4 = 0; j4 < A.tw4p; ++j4) {
                const uint32_t w1 = Lp[j4 + 1];
                const uint32_t o = A.ones_words[j4];
                s0 = dp4a_u(w0, o, s0);
                s1 = dp4a_u(__funnelshift_r(w0, w1, 8), o, s1);
                s2 = dp4a_u(__funnelshift_r(w0, w1, 16), o, s2);
                s3 = dp4a_u(__funnelshift_r(w0, w1, 24), o, s3);
                w0 = w1;
            }
            uint16_t *hp = Hs + (size_t)row * (XT * 4) + g * 4;
            hp[0] = (uint16_t)s0; hp[1] = (uint16_t)s1; hp[2] = (uint16_t)s2; hp[3] = (uint16_t)s3;
        }
    }

    // ---- 3. the correlation: 4 x RY register tile per thread ---------------------------
    int acc[RY][4];
#pragma unroll
    for (int r = 0; r < RY; ++r)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[r][k] = 0;

    const uint32_t *Lbase = Ls + (size_t)(ty * RY) * A.pitch_w + tx;
    const int npairs = A.tw4p >> 1;
    for (int jp = 0; jp < npairs; ++jp) {
        const uint2 *Tp = reinterpret_cast<const uint2 *>(Ts) + (size_t)jp * A.steps;
        const uint32_t *Lp = Lbase + 2 * jp;
        uint32_t wa[RY], wb[RY];        // sliding windows of template words (two columns groups)
#pragma unroll
        for (int r = 0; r < RY; ++r) { wa[r] = 0; wb[r] = 0; }
        for (int s0 = 0; s0 < A.steps; s0 += RY) {
#pragma unroll
            for (int u = 0; u < RY; ++u) {
                const int s = s0 + u;                   // lightness row y0 + s, template row s
                const uint32_t *Lr = Lp + (size_t)s * A.pitch_w;
                const uint32_t l0 = Lr[0], l1 = Lr[1], l2 = Lr[2];
                const uint2 t = Tp[s];
                wa[u] = t.x; wb[u] = t.y;               // template row s (zero for s >= th)
                const uint32_t a0 = l0, a1 = __funnelshift_r(l0, l1, 8),
                               a2 = __funnelshift_r(l0, l1, 16), a3 = __funnelshift_r(l0, l1, 24);
                const uint32_t b0 = l1, b1 = __funnelshift_r(l1, l2, 8),
                               b2 = __funnelshift_r(l1, l2, 16), b3 = __funnelshift_r(l1, l2, 24);
#pragma unroll
                for (int r = 0; r < RY; ++r) {
                    // result row y0 + r sees template row i = s - r, held in slot (u - r) mod RY
                    const uint32_t ta = wa[(u - r + RY) % RY], tb = wb[(u - r + RY) % RY];
                    acc[r][0] = dp4a_u(a0, ta, acc[r][0]);
                    acc[r][1] = dp4a_u(a1, ta, acc[r][1]);
                    acc[r][2] = dp4a_u(a2, ta, acc[r][2]);
                    acc[r][3] = dp4a_u(a3, ta, acc[r][3]);
                    acc[r][0] = dp4a_u(b0, tb, acc[r][0]);
                    acc[r][1] = dp4a_u(b1, tb, acc[r][1]);
                    acc[r][2] = dp4a_u(b2, tb, acc[r][2]);
                    acc[r][3] = dp4a_u(b3, tb, acc[r][3]);
                }
            }
        }
    }
    __syncthreads();        // Hs complete (written by other threads in step 2)

    // ---- 4. window sums for this thread's results, score, arg-max -----------------------
    u64 best = 0;
    {
        int win[4] = {0, 0, 0, 0};
        const uint16_t *hp = Hs + (size_t)(ty * RY) * (XT * 4) + tx * 4;
        for (int i = 0; i < A.tmpl_h; ++i) {
            const uint2 v = *reinterpret_cast<const uint2 *>(hp + (size_t)i * (XT * 4));
            win[0] += v.x & 0xffff; win[1] += v.x >> 16; win[2] += v.y & 0xffff; win[3] += v.y >> 16;
        }
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            const int y = tile_y0 + ty * RY + r;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int x = tile_x0 + tx * 4 + k;
                if (y < A.geom.res_h && x < A.geom.res_w) {
                    const float v = melf_ccoeff_value(A.n_t, A.sum_t, acc[r][k], win[k]);
                    const uint32_t idx = (uint32_t)y * (uint32_t)A.geom.res_w + (uint32_t)x;
                    if (A.result)
                        A.result[(size_t)f * A.geom.res_w * A.geom.res_h + idx] = v;
                    const u64 key = melf_score_key(v, idx);
                    best = key > best ? key : best;
                }
            }
            // slide the window down one row
            const uint2 top = *reinterpret_cast<const uint2 *>(hp + (size_t)r * (XT * 4));
            const uint2 bot = *reinterpret_cast<const uint2 *>(hp + (size_t)(r + A.tmpl_h) * (XT * 4));
            win[0] += (int)(bot.x & 0xffff) - (int)(top.x & 0xffff);
            win[1] += (int)(bot.x >> 16) - (int)(top.x >> 16);
            win[2] += (int)(bot.y & 0xffff) - (int)(top.y & 0xffff);
            win[3] += (int)(bot.y >> 16) - (int)(top.y >> 16);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other > best ? other : best;
    }
    __shared__ u64 warp_best[32];
    const int warp = tid >> 5, lane = tid & 31;
    if (lane == 0) warp_best[warp] = best;
    __syncthreads();
    if (tid == 0) {
        const int nwarps = (nthreads + 31) >> 5;
        for (int w = 1; w < nwarps; ++w) best = warp_best[w] > best ? warp_best[w] : best;
        if (best) atomicMax(A.match_key + f, best);
    }
}

// unpack the per-frame key into (x, y) and the score: the tail of match_template
__global__ void melf_match_unpack_kernel(const u64 *key, int n, int res_w, int32_t *xy, float *val)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 k = key[i];
    const uint32_t idx = melf_key_index(k);
    xy[2 * i] = (int32_t)(idx % (uint32_t)res_w);
    xy[2 * i + 1] = (int32_t)(idx / (uint32_t)res_w);
    val[i] = melf_key_score(k);
}

size_t melf_match_smem_bytes(const MelfMatchArgs &a)
{
    return (size_t)a.tile_rows * a.pitch_w * 4 + (size_t)a.tw4p * a.steps * 4 +
           (size_t)a.tile_rows * (XT * 4) * 2;
}

// Fill the tiling fields of args for a template of th x tw; returns the grid.
dim3 melf_match_plan(MelfMatchArgs &a, int n_frames)
{
    const int rw = a.geom.res_w, rh = a.geom.res_h;
    a.tw4p = (((a.tmpl_w + 3) / 4) + 1) & ~1;
    a.steps = ((a.tmpl_h + RY - 1 + RY - 1) / RY) * RY;
    // rows of results per CTA: split the result height evenly over the fewest tiles whose
    // shared memory still lets 2 CTAs share an SM (<= ~100 KB each)
    const int groups = (rh + RY - 1) / RY;
    int ntiles = 1;
    for (;; ++ntiles) {
        a.nyt = (groups + ntiles - 1) / ntiles;
        a.tile_rows = a.nyt * RY + a.steps;
        // 3 words are read per step starting at word tx + 2*jp: up to 15 + tw4p-2 + 2
        a.pitch_w = XT + a.tw4p;
        while ((a.pitch_w & 3) != 2) ++a.pitch_w;      // spreads the two y-groups of a warp over banks
        if ((a.nyt <= 32) && melf_match_smem_bytes(a) <= 100 * 1024) break;
        if (a.nyt == 1) break;
    }
    return dim3((rw + XT * 4 - 1) / (XT * 4), ntiles, n_frames);
}

cudaError_t melf_launch_match(const MelfMatchArgs &args, dim3 grid, cudaStream_t stream)
{
    const size_t smem = melf_match_smem_bytes(args);
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(melf_match_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    melf_match_kernel<<<grid, dim3(XT, args.nyt), smem, stream>>>(args);
    return cudaGetLastError();
}

cudaError_t melf_launch_match_unpack(const u64 *key, int n, int res_w, int32_t *xy, float *val,
                                     cudaStream_t stream)
{
    melf_match_unpack_kernel<<<(n + 255) / 256, 256, 0, stream>>>(key, n, res_w, xy, val);
    return cudaGetLastError();
}
