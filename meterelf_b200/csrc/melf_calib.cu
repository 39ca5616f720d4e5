// Offline calibration support: the average meter image.
//
// Replaces get_average_meter_image (_calibration.py:63-66) = for every file
// ImageFile.get_bgr_image_t (_image.py:34-44: cv2.warpAffine by the integer translation that
// moves the matched dials rect to a fixed anchor, zero border), normalize_image
// (_utils.py:64-65: float64 / 255.0), the running mean of calculate_average_of_norm_images /
// _image_avg_reducer (_utils.py:72-88: new = prev * ((n - 1) / n) + img / n, n = 2, 3, ...) and
// denormalize_image (_utils.py:68-69: uint8(avg * 255.0 + 0.5), truncating).
//
// One thread per output byte; the frames are visited in order, so the float64 recurrence is the
// reference's operation for operation (no FMA contraction: explicit _rn intrinsics).  An integer
// translation under cv2's bilinear warp is an exact pixel copy.
#include "melf_common.cuh"

namespace {

__global__ void __launch_bounds__(256)
melf_average_kernel(const uint8_t *bgr, int n, int h, int w, size_t row_stride, size_t frame_stride,
                    const int32_t *xy, int anchor_x, int anchor_y, uint8_t *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int row_bytes = 3 * w;
    if (i >= h * row_bytes) return;
    const int y = i / row_bytes, xb = i - y * row_bytes;          // xb = 3 * x + channel
    double avg = 0.0;
    for (int f = 0; f < n; ++f) {
        // dst(x, y) = src(x - tx, y - ty), (tx, ty) = anchor - match top-left (_image.py:39-44)
        const int tx = anchor_x - xy[2 * f], ty = anchor_y - xy[2 * f + 1];
        const int sy = y - ty, sxb = xb - 3 * tx;
        int v = 0;
        if (sy >= 0 && sy < h && sxb >= 0 && sxb < row_bytes)
            v = bgr[(size_t)f * frame_stride + (size_t)sy * row_stride + sxb];
        const double norm = __ddiv_rn((double)v, 255.0);
        if (f == 0) {
            avg = norm;
        } else {
            const double cnt = (double)(f + 1);
            avg = __dadd_rn(__dmul_rn(avg, __ddiv_rn(cnt - 1.0, cnt)), __ddiv_rn(norm, cnt));
        }
    }
    out[i] = (uint8_t)(int)__dadd_rn(__dmul_rn(avg, 255.0), 0.5);
}

}  // namespace

extern "C" int melf_average_translated(int device, const uint8_t *d_bgr, int n, int h, int w,
                                       size_t row_stride, size_t frame_stride, const int32_t *d_match_xy,
                                       int anchor_x, int anchor_y, uint8_t *d_out, void *stream)
{
    if (n < 1 || h < 1 || w < 1 || !d_bgr || !d_match_xy || !d_out || row_stride < (size_t)3 * w)
        return MELF_ERR_INVALID;
    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) return MELF_ERR_CUDA;
    const int total = h * w * 3;
    melf_average_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        d_bgr, n, h, w, row_stride, frame_stride, d_match_xy, anchor_x, anchor_y, d_out);
    const cudaError_t e = cudaGetLastError();
    if (prev != device) cudaSetDevice(prev);      // leave the caller's current device as it was
    return e == cudaSuccess ? MELF_OK : MELF_ERR_CUDA;
}
