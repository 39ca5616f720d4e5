// Probe: does a tiled tensor copy (cp.async.bulk.tensor.3d) of u8 data accept a start coordinate that is not a
// multiple of 16 bytes, and which box widths does it take?   nvcc -arch=sm_100a -o tmap_probe tmap_probe.cu
//     ./tmap_probe <c0> <box_w> [rank3=1]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

__global__ void probe(const __grid_constant__ CUtensorMap map, int c0, int c1, int c2, int box_bytes, uint8_t *out)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ unsigned long long bar;
    const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(&bar);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar_s));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_s), "r"(box_bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
                     " [%0], [%1, {%2, %3, %4}], [%5];"
                     :: "r"(dst), "l"(reinterpret_cast<uint64_t>(&map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar_s) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p;}"
                     : "=r"(done) : "r"(bar_s) : "memory");
    for (int i = threadIdx.x; i < box_bytes; i += blockDim.x) out[i] = smem[i];
}

int main(int argc, char **argv)
{
    const int c0 = argc > 1 ? atoi(argv[1]) : 0, box_w = argc > 2 ? atoi(argv[2]) : 208;
    const int W = 1440, H = 640, N = 4, rows = 8;
    std::vector<uint8_t> h((size_t)W * H * N);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *d_out;
    cudaMalloc(&d, h.size());
    cudaMalloc(&d_out, 65536);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                 const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (!fn) { printf("no entry point\n"); return 2; }
    alignas(64) CUtensorMap map;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    const cuuint64_t strides[2] = {(cuuint64_t)W, (cuuint64_t)W * H};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)rows, 1};
    const cuuint32_t elem[3] = {1, 1, 1};
    CUresult r = reinterpret_cast<EncodeFn>(fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, elem,
                                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("c0=%d box_w=%d encode=%d\n", c0, box_w, (int)r);
    if (r != CUDA_SUCCESS) return 3;
    const int c1 = 637, c2 = 2, box_bytes = box_w * rows;       // the last rows reach past the frame: zeros expected
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    probe<<<1, 128, 65536>>>(map, c0, c1, c2, box_bytes, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 4;
    std::vector<uint8_t> got(box_bytes);
    cudaMemcpy(got.data(), d_out, box_bytes, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < rows; ++rr)
        for (int x = 0; x < box_w; ++x) {
            const int gx = c0 + x, gy = c1 + rr;
            const uint8_t want = (gx < W && gy < H && gx >= 0) ? h[((size_t)c2 * H + gy) * W + gx] : 0;
            bad += got[rr * box_w + x] != want;
        }
    printf("mismatching bytes: %d of %d\n", bad, box_bytes);
    return bad ? 5 : 0;
}
