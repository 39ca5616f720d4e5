// Kernel argument blocks shared between melf_api.cu and the kernel translation units.
#pragma once
#include "melf_common.cuh"

// per-frame hand-over between the dial warps of melf_dials_kernel (they sit in different CTAs):
// each warp deposits its dial, the last one to arrive (count) assembles the frame result and
// resets count for the next launch
struct MelfDialSlot {
    double pos[MELF_MAX_DIALS];
    int32_t flag[MELF_MAX_DIALS];   // 0 read, 1 no contours, 2 unreadable
    uint32_t count;
    uint32_t pad;
};

struct MelfDialsArgs {
    const uint8_t *bgr;
    const uint8_t *bgr_end;      // one past the last readable byte of the frame memory
    size_t row_stride, frame_stride;
    MelfGeom geom;
    const MelfDialDev *dials;
    const uint16_t *need_pos;    // MelfDialDev::need_off slices
    const uint16_t *ring_pos;
    const double *ring_angle;
    const double *ring_sqd;
    const u64 *match_key;
    melf_result *out;
    MelfDialSlot *slots;         // [n_frames], count zero on entry
    int32_t n_frames;
    melf_taps taps;
    double threshold;
    int32_t n_dials, tmpl_w, tmpl_h, hue_shift, s_tail;
    int32_t value_order[4];
    int32_t pix_pitch, pix_bytes; // per-dial shared-memory window: row pitch, total bytes (multiples of 16)
    long long *timeline;         // optional debug: clock64 stamps of (frame 0 .. 7, dial 0), 8 per frame
};

struct MelfMatchArgs {
    const uint8_t *img;          // BGR (channels = 3) or single channel frames
    size_t row_stride, frame_stride;
    int32_t channels;
    MelfGeom geom;
    const uint32_t *tmpl_words;  // [tw4p/2][steps][2]: template, 4 columns per word
    const uint32_t *ones_words;  // [tw4p]: 1 per valid template column
    int32_t tmpl_w, tmpl_h;
    int32_t tw4p;                // template width in words, padded to even
    int32_t steps;               // roundup(th + RY - 1, RY)
    int32_t nyt;                 // y-threads per CTA
    int32_t pitch_w;             // lightness tile row pitch in words
    int32_t tile_rows;           // rows of the lightness tile = nyt*RY + steps - RY ... (+slack)
    long long sum_t, n_t;
    u64 *match_key;              // [n] atomicMax slot per frame (zeroed before launch)
    float *result;               // optional [n][res_h][res_w]
};

// tensor-core TM_CCOEFF (melf_match_tc.cu)
struct MelfMatchTcArgs {
    const uint8_t *img;          // BGR (channels = 3) or single channel frames
    const uint8_t *img_limit;    // 16-byte aligned end of the readable image memory
    size_t row_stride, frame_stride;
    int32_t channels;
    MelfGeom geom;
    const uint8_t *toeplitz;     // [n_pairs][16 strips][128][16 B] Toeplitz operands of the template
    int32_t tmpl_w, tmpl_h, n_pairs;
    int32_t kpp;                 // K-steps (of 32) of a pair that can hold a template byte: ceil((63 + tmpl_w) / 32) <= 8
    int32_t n_tx, n_ty, chunk_h; // work items per frame: 64-column tiles x row chunks
    long long n_items;
    long long sum_t, n_t;
    u64 *match_key;              // [n] atomicMax slot per frame (zeroed before launch)
    float *result;               // optional [n][res_h][res_w]
    uint32_t ones_words[64];     // byte j = 1 for j < tmpl_w
    // scratch written by the preparation kernel, read by the tensor-core kernel; one launch
    // pair covers the items [item0, item0 + max_items) (max_items even)
    uint8_t *slabs;              // [max_items][64 KB] lightness slabs in the UMMA strip layout
    uint32_t *wsum;              // [max_items][64][137] window sums, x-major
    long long item0, max_items;
    long long *timeline;         // optional debug: clock64 stamps of CTA 0's first 8 passes, 8 per pass
    int32_t img_w, img_h;        // whole frames as the caller holds them (tensor map of the batch)
    long long img_n;
    int32_t use_tmap;            // set by the launcher: frame bytes staged by tiled tensor copies
    int32_t debug_skip;          // measurement only (MELF_TC_DEBUG_SKIP, results become wrong): bit 0 no staging after
                                 // the prologue, bit 1 no window sums, bit 2 no drain, bit 3 no MMAs,
                                 // bit 4 staging without its global loads, bit 5 staging without the conversion
};

// Small result maps (sample set 2: 33 x 17 positions): result positions as the MMA's M, frames as its N
// (melf_match_tc.cu, last section).
struct MelfSmallArgs {
    const uint8_t *img;          // BGR (channels = 3) or single channel frames
    const uint8_t *img_limit;    // 16-byte aligned end of the readable image memory
    size_t row_stride, frame_stride;
    int32_t channels;
    MelfGeom geom;
    int32_t n_frames;
    int32_t tmpl_w, tmpl_h;
    int32_t n_pos, n_mt;         // res_w * res_h result positions (+ res_w box-sum rows) in n_mt tiles of 128
    int32_t cwp;                 // crop width padded to 32: K index of crop pixel (r, c) = r * cwp + c
    int32_t n_kstages;           // ceil(crop_h * cwp / 64)
    int32_t mt_k0[5], mt_k1[5];  // K-stage range in which tile mt has a non-zero template byte
    const uint8_t *tpos;         // template per position: [n_kstages][n_mt][2 K-steps][2 strips][128 positions][16 B]
    uint32_t *wsum;              // edge sums E(y, x): [group][n_mt * 128 positions][32 frames]
    long long sum_t, n_t;
    u64 *match_key;              // [n_frames]
};

struct MelfConvertArgs {
    const uint8_t *bgr;
    size_t row_stride, frame_stride;
    int32_t n, h, w;
    int32_t hue_shift, s_tail;
    uint8_t *hls;      // [n][h][w][3] or NULL
    uint8_t *light;    // [n][h][w]    or NULL
    uint8_t *mask;     // [n][h][w]    or NULL
    uint32_t lo, hi;   // packed H | L<<8 | S<<16 bounds (inclusive)
    // exact 32-bit division by multiplication (Granlund-Montgomery) for the flat index
    uint32_t gdiv[3];  // by w / 16: multiplier, shift 1, shift 2
    uint32_t hdiv[3];  // by h
};

__host__ __device__ __forceinline__ uint32_t melf_fastdiv(uint32_t n, const uint32_t (&d)[3])
{
#ifdef __CUDA_ARCH__
    const uint32_t t = __umulhi(d[0], n);
#else
    const uint32_t t = (uint32_t)(((unsigned long long)d[0] * n) >> 32);
#endif
    return (t + ((n - t) >> d[1])) >> d[2];
}
