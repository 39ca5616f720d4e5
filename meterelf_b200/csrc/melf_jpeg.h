// Internal interface of the JPEG ingest stage (melf_jpeg.cu) towards melf_api.cu.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>
#include <string>

struct MelfJpegState;                    // scratch buffers, grown on demand, one per stream slot

struct MelfJpegBatch {                   // host-side view of one call
    const uint8_t *jpeg;                 // HOST: concatenated streams
    const uint64_t *offsets;             // HOST: n + 1 offsets into jpeg
    int n;
};

// Host-only header probe of one stream: 0 or MELF_ERR_* with a text in *err.
int melf_jpeg_probe(const uint8_t *jpeg, size_t len, int *h, int *w, int *ncomp, int *hs, int *vs,
                    std::string *err);

// Decodes streams [first, first+count) of `batch` into BGR pixels of the rectangle
// [x0, x0+cw) x [y0, y0+ch) of every frame: frame f of the call at d_out + f*frame_stride, rows
// row_stride apart.  Every stream must decode to h x w pixels.  Asynchronous on `stream`; the
// host staging buffer of `*state` is reused by the next call on the same state, which therefore
// waits for this call's upload.  Returns 0 or MELF_ERR_*; *launches counts kernel launches.
int melf_jpeg_decode_range(MelfJpegState **state, const MelfJpegBatch &batch, int first, int count,
                           int h, int w, int x0, int y0, int cw, int ch, uint8_t *d_out,
                           size_t row_stride, size_t frame_stride, cudaStream_t stream,
                           std::string *err, uint64_t *launches, cudaEvent_t (*spans)[2] = nullptr,
                           cudaStream_t copy_stream = nullptr);
// spans (optional): 5 event pairs recorded around upload, destuff, Huffman, IDCT, colour.
// copy_stream (optional): the stream the host-to-device copy is issued on (`stream` waits for it).

void melf_jpeg_state_free(MelfJpegState *state);
