// PCIe probe (measurement tool, not product): H2D rate of (a) one flat copy, (b) the
// meter_rect crop as cudaMemcpy3DAsync (750-byte rows out of 1440-byte frame rows),
// (c) a gather kernel reading mapped pinned host memory directly.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void gather_rows(const uint8_t *src, uint8_t *dst, int n, int h, int w3, int x0b, int y0,
                            int ch, int cwb, size_t pitch)
{
    // one warp per crop row: 16-byte aligned loads covering [x0b, x0b+cwb)
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nw = (gridDim.x * blockDim.x) >> 5;
    for (long long r = warp; r < (long long)n * ch; r += nw) {
        const int f = (int)(r / ch), y = (int)(r % ch);
        const uint8_t *s = src + ((size_t)f * h + y0 + y) * w3 + x0b;
        const uintptr_t a0 = (uintptr_t)s & ~(uintptr_t)15;
        const int nvec = (int)(((uintptr_t)s + cwb + 15 - a0) >> 4);
        uint4 *d = (uint4 *)(dst + (size_t)r * pitch);
        for (int v = lane; v < nvec; v += 32) d[v] = ((const uint4 *)a0)[v];
    }
}

int main()
{
    const int n = 2048, h = 640, w = 480, x0 = 100, y0 = 150, cw = 250, ch = 250;
    const size_t fb = (size_t)h * w * 3, total = fb * n, crop = (size_t)cw * 3 * ch * n;
    uint8_t *hbuf, *dbuf, *dcrop;
    CK(cudaHostAlloc(&hbuf, total, cudaHostAllocMapped));
    memset(hbuf, 1, total);
    CK(cudaMalloc(&dbuf, total));
    const size_t pitch = 768;
    CK(cudaMalloc(&dcrop, pitch * ch * n + 4096));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0); CK(cudaMemcpyAsync(dbuf, hbuf, total, cudaMemcpyHostToDevice)); cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
        printf("flat H2D      %.1f MB: %.2f GB/s\n", total / 1e6, total / ms / 1e6);
        cudaEventRecord(e0); CK(cudaMemcpyAsync(dbuf, hbuf, crop, cudaMemcpyHostToDevice)); cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
        printf("flat H2D      %.1f MB: %.2f GB/s\n", crop / 1e6, crop / ms / 1e6);
        cudaMemcpy3DParms p{};
        p.srcPtr = make_cudaPitchedPtr(hbuf, (size_t)w * 3, (size_t)w * 3, h);
        p.srcPos = make_cudaPos((size_t)x0 * 3, y0, 0);
        p.dstPtr = make_cudaPitchedPtr(dcrop, pitch, (size_t)cw * 3, ch);
        p.extent = make_cudaExtent((size_t)cw * 3, ch, n);
        p.kind = cudaMemcpyHostToDevice;
        cudaEventRecord(e0); CK(cudaMemcpy3DAsync(&p, 0)); cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
        printf("3D crop H2D   %.1f MB: %.2f GB/s (useful bytes)\n", crop / 1e6, crop / ms / 1e6);
        // rows only (full-width rows y0..y0+ch): 2D copy, contiguous 360 KB per frame
        cudaEventRecord(e0);
        CK(cudaMemcpy2DAsync(dbuf, (size_t)ch * w * 3, hbuf + (size_t)y0 * w * 3, fb, (size_t)ch * w * 3, n, cudaMemcpyHostToDevice));
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
        printf("row-band H2D  %.1f MB: %.2f GB/s raw, %.2f GB/s useful\n", (double)ch * w * 3 * n / 1e6,
               (double)ch * w * 3 * n / ms / 1e6, crop / ms / 1e6);
        uint8_t *dmap; CK(cudaHostGetDevicePointer(&dmap, hbuf, 0));
        for (int blocks = 148; blocks <= 148 * 16; blocks *= 4) {
            cudaEventRecord(e0);
            gather_rows<<<blocks, 256>>>(dmap, dcrop, n, h, w * 3, x0 * 3, y0, ch, cw * 3, pitch);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
            printf("zero-copy gather, %d blocks: %.2f GB/s (useful bytes)\n", blocks, crop / ms / 1e6);
        }
    }
    return 0;
}
