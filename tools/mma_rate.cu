// tcgen05.mma.kind::i8 issue-rate probe: cycles per 128 x N x 32 instruction as a function of N
// (A and B from shared memory, K-major no-swizzle as in melf_match_tc.cu).  Used to size the MMA N of the
// TM_CCOEFF kernel; output committed under profiles/.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mma_rate tools/mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ u64 umma_desc(uint32_t addr, uint32_t lbo_bytes)
{
    return (u64)((addr >> 4) & 0x3fffu) | ((u64)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((u64)((128u >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
__host__ __device__ constexpr uint32_t make_idesc(uint32_t m, uint32_t n)
{
    return (2u << 4) | ((n >> 3) << 17) | ((m >> 4) << 24);
}
extern __shared__ __align__(1024) uint8_t smem[];

__global__ void __launch_bounds__(128, 1) rate_kernel(int n, int m, int reps, int a_tmem, int sh, long long *out)
{
    __shared__ unsigned long long bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x;
    for (int i = tid; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0x01010101u * (i & 3);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    if (warp == 0) {
        uint32_t lead;
        asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(lead));
        const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 65536);
        const uint32_t idesc = make_idesc((uint32_t)m, (uint32_t)n);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                // `run` consecutive instructions accumulate into one accumulator, then the other (0 = one only)
                const uint32_t d = tmem + ((((r * 8 + ks) >> sh) & 1) ? 192u : 0u);
                if (lead) {
                const u64 ad = umma_desc(a_addr + ks * 4096, 2048), bd = umma_desc(b_addr + ks * 8448 + (r & 63) * 32, 4224);
                if (a_tmem)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                                 :: "r"(d), "r"(tmem + 448 + 8 * ks), "l"(bd), "r"(idesc), "r"(1) : "memory");
                else
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                                 :: "r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1) : "memory");
                }
            }
        }
        if (lead) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
        uint32_t done;
        do {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
        } while (!done);
        const long long t1 = clock64();
        if (blockIdx.x == 0 && lead) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

int main()
{
    long long *d, h;
    cudaMalloc(&d, 8);
    cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int reps = 256;
    printf("# M N a_tmem run cycles_per_mma MAC_per_clk (all 148 SMs busy; run = log2(consecutive MMAs per accumulator), 31 = single accumulator; one elected lane of a converged warp issues)\n");
    const int runs[4] = {31, 0, 1, 3};     // log2 of the run length; 31 = single accumulator
    for (int m = 128; m >= 64; m -= 64)
        for (int a_tmem = 0; a_tmem < 2; ++a_tmem)
          for (int ri = 0; ri < 4; ++ri)
            for (int n = 16; n <= 192; n += 16) {
                const int run = runs[ri];
                if (m == 64 && ri) continue;
                rate_kernel<<<148, 128, 200 * 1024>>>(n, m, reps, a_tmem, run, d);
                rate_kernel<<<148, 128, 200 * 1024>>>(n, m, reps, a_tmem, run, d);
                if (cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                const double cyc = (double)h / (reps * 8);
                printf("%d %d %d %d %.1f %.0f\n", m, n, a_tmem, run, cyc, (double)m * n * 32 / cyc);
            }
    return 0;
}
