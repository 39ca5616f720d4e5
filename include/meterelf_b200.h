/*
 * meterelf_b200.h -- C ABI of the B200-native dial-reading hot path.
 *
 * The reference (suutari/meterelf) is pure Python and has no FFI of its own; the seam
 * is cut at the functions below.  Each entry point names the
 * reference interface it replaces.  Plain pointers and sizes only, no exceptions cross
 * this boundary: API/CUDA failures are integer return codes (0 = MELF_OK) with a text
 * in melf_last_error(); per-frame data outcomes are melf_result.status.
 *
 * Threading: one melf_ctx per GPU; a ctx is not thread-safe; independent ctxs may be
 * used concurrently.  The caller owns every buffer it passes; a ctx owns its device
 * scratch.  Pointers named d_* must be device pointers on the ctx's device.
 */
#ifndef METERELF_B200_H
#define METERELF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MELF_ABI_VERSION 1
#define MELF_MAX_DIALS 8
/* per-dial working window (disc bounding box plus a 2-pixel closing halo) is held as
 * 64 rows of 64-bit words, so a dial's mask may span at most 60 x 60 pixels */
#define MELF_DIAL_WINDOW 64
#define MELF_DIAL_MAX_EXTENT 60

/* return codes */
enum {
    MELF_OK = 0,
    MELF_ERR_INVALID = 1,       /* bad argument / params                               */
    MELF_ERR_CUDA = 2,          /* a CUDA call failed; see melf_last_error             */
    MELF_ERR_UNSUPPORTED = 3,   /* valid for the reference but outside this build's limits */
    MELF_ERR_TOO_SMALL = 4      /* meter_rect crop smaller than the dials template:
                                   cv2.matchTemplate would raise cv2.error there       */
};

/* melf_result.status: which exception the reference would raise for the frame */
enum {
    MELF_ST_OK = 0,
    MELF_ST_DIALS_NOT_FOUND = 1,     /* DialsNotFoundError,          _image.py:62-64     */
    MELF_ST_NO_CONTOURS = 2,         /* NeedleContoursNotFoundError, _reading.py:137-138 */
    MELF_ST_ANGLE_UNDETERMINED = 3   /* DialAngleDeterminingError,   _reading.py:98-106  */
};

/* One dial = one entry of params.yml needle_data (_params.py:71-81) plus its DialData
 * masks (_types.py:17-21, _dial_data.py:22-55). */
typedef struct melf_dial {
    double center_x, center_y;     /* float centre, dials-template coordinates          */
    double angle_of_zero;          /* degrees                                            */
    int32_t range_h, range_l, range_s; /* color_range                                    */
    int32_t negative_momentum;     /* 0/1                                                */
    const uint8_t *mask;           /* tmpl_h*tmpl_w u8, nonzero = inside the dial disc   */
    const uint8_t *circle_mask;    /* tmpl_h*tmpl_w u8, nonzero = needle-tip annulus     */
} melf_dial;

/* Everything the path needs from Params (_params.py:30-64). Host pointers; copied. */
typedef struct melf_params {
    uint32_t struct_size;          /* sizeof(melf_params), ABI check                     */
    int32_t n_dials;               /* 0..MELF_MAX_DIALS, YAML order = evaluation order;
                                      0 = locate-only ctx (melf_match_template only)     */
    int32_t tmpl_w, tmpl_h;        /* dials_template_size                                */
    const uint8_t *tmpl;           /* tmpl_h*tmpl_w grayscale template (_image.py:72-81) */
    int32_t rect_x0, rect_y0, rect_x1, rect_y1; /* meter_rect; clipped to the frame like
                                      the reference's numpy slice (_utils.py:52-54)      */
    int32_t hue_shift;             /* added to H modulo 256 (_utils.py:102)              */
    int32_t s_tail;                /* 1: reproduce OpenCV's scalar-tail S rounding in the
                                      last (w mod 8) columns of each 256-pixel block     */
    double match_threshold;        /* dials_template_match_threshold                     */
    int32_t value_order[4];        /* dial indices sorted by dial NAME (_reading.py:171) */
    melf_dial dials[MELF_MAX_DIALS];
} melf_params;

/* Per-frame outcome = return value / exception of get_meter_value (_reading.py:19-115). */
typedef struct melf_result {
    int32_t status;                /* MELF_ST_*                                          */
    int32_t match_x, match_y;      /* TemplateMatchResult.rect.top_left (crop coords)    */
    float max_val;                 /* TemplateMatchResult.max_val                        */
    int32_t failed_dial;           /* MELF_ST_NO_CONTOURS: first dial without contours   */
    uint32_t unreadable_mask;      /* MELF_ST_ANGLE_UNDETERMINED: bit d = dial d         */
    double dial_pos[MELF_MAX_DIALS]; /* dial positions 0..10, NaN where not available    */
    double value;                  /* meter value, NaN if absent                         */
} melf_result;

/* Optional per-(frame, dial) debug taps (device buffers, any may be NULL).  Row r of a
 * tap window is a 64-bit word, bit c = pixel (win_x0 + c, win_y0 + r) of the dials
 * rect, where (win_x0, win_y0) = melf_dial_window(). */
typedef struct melf_taps {
    uint64_t *d_mask_orig;         /* [n][n_dials][64] inRange result                    */
    uint64_t *d_mask_de;           /* [n][n_dials][64] after 3x3 dilate + erode          */
    uint64_t *d_needle_mask;       /* [n][n_dials][64] selected needle mask              */
    int32_t *d_dial_color;         /* [n][n_dials][3]  get_dial_color                    */
    int32_t *d_counts;             /* [n][n_dials][4]  needle points, kept ring points,
                                      2*area of the selected contour (-1 none), components */
    double *d_momentum;            /* [n][n_dials][3]  momentum x, y, angle (NaN = None) */
} melf_taps;

typedef struct melf_ctx melf_ctx;

/* ---- context ------------------------------------------------------------------- */
/* Replaces the per-params caches of the reference (_image.py:69-81 template,
 * _dial_data.py:11-19 masks): uploads template + masks and precomputes the per-dial
 * static tables. */
int melf_create(melf_ctx **out, int device, const melf_params *params);
void melf_destroy(melf_ctx *ctx);
/* Text of the last failure on ctx (ctx may be NULL: last melf_create failure). */
const char *melf_last_error(const melf_ctx *ctx);
int melf_abi_version(void);
/* Top-left corner of dial d's 64x64 tap window in dials-rect coordinates. */
int melf_dial_window(const melf_ctx *ctx, int dial, int32_t *win_x0, int32_t *win_y0);

/* ---- measurement ----------------------------------------------------------------- */
/* While enabled, every kernel launch of the path on this ctx is bracketed by CUDA events on
 * its own stream.  melf_profile_read synchronises the device, adds up the elapsed
 * milliseconds and launch counts per kernel kind since the last read and clears them.
 * For bench.py's roofline numbers (no reference counterpart). */
#define MELF_PROF_MATCH_PREP 0      /* melf_match_prep_kernel: lightness slabs + window sums */
#define MELF_PROF_MATCH_TC   1      /* melf_match_tc_kernel: tcgen05 correlation + arg-max    */
#define MELF_PROF_DIALS      2      /* melf_dials_kernel                                      */
#define MELF_PROF_MATCH_DP4A 3      /* melf_match_kernel: dp4a correlation (large templates)  */
#define MELF_PROF_JPEG_UPLOAD  4    /* melf_decode_jpeg: H2D copy of the streams + coefficient clear */
#define MELF_PROF_JPEG_DESTUFF 5    /* melf_jpeg_destuff_kernel                               */
#define MELF_PROF_JPEG_HUFFMAN 6    /* melf_jpeg_huffman_kernel                               */
#define MELF_PROF_JPEG_IDCT    7    /* melf_jpeg_idct_kernel                                  */
#define MELF_PROF_JPEG_COLOR   8    /* melf_jpeg_color_kernel                                 */
#define MELF_PROF_KINDS      9
int melf_profile_enable(melf_ctx *ctx, int on);
/* Name of the locate kernel the last call on this ctx took (size dispatch: Toeplitz tcgen05 kernel, small-map
 * GEMM, dp4a kernel for templates the tensor kernels do not fit); for measurement records. */
const char *melf_last_locate(const melf_ctx *ctx);
int melf_profile_read(melf_ctx *ctx, float *ms /*[MELF_PROF_KINDS]*/, int32_t *launches /*[MELF_PROF_KINDS]*/);

/* ---- the hot path -------------------------------------------------------------- */
/* Batched ImageFile.get_dials_hls + get_meter_value (_image.py:23-27, _reading.py:19-115)
 * over n BGR frames of h x w pixels already resident on the device.  Frame f starts at
 * d_bgr + f*frame_stride, row y at + y*row_stride (bytes).  Asynchronous on `stream`
 * (a cudaStream_t, may be NULL); d_out receives n melf_result records (device memory).
 * taps may be NULL. */
int melf_read_batch_async(melf_ctx *ctx, const uint8_t *d_bgr, int n, int h, int w,
                          size_t row_stride, size_t frame_stride,
                          melf_result *d_out, const melf_taps *taps, void *stream);

/* Same, with HOST frames and HOST results: copies the meter_rect rows host->device in
 * chunks on two streams (copy of chunk k+1 overlaps the kernels of chunk k), runs the
 * path and copies the results back.  Synchronous.  bgr should be pinned for full PCIe
 * speed but pageable memory works. */
int melf_read_batch(melf_ctx *ctx, const uint8_t *bgr, int n, int h, int w,
                    size_t row_stride, size_t frame_stride, melf_result *out);

/* ---- stage operators (drop-ins for the _utils.py primitives) --------------------- */
/* convert_to_hls (_utils.py:100-102) fused with get_mask_by_color (_utils.py:113-119):
 * one pass over n frames of h x w BGR pixels.  Any of d_hls (n*h*w*3, interleaved
 * H,L,S), d_light (n*h*w, the L plane that _image.py:59 splits off) and d_mask (n*h*w,
 * 255 where lo <= HLS <= hi per channel) may be NULL.  lo/hi: 3 bytes each, host. */
int melf_convert_hls(int device, const uint8_t *d_bgr, int n, int h, int w,
                     size_t row_stride, size_t frame_stride, int hue_shift, int s_tail,
                     uint8_t *d_hls, uint8_t *d_light,
                     const uint8_t *lo, const uint8_t *hi, uint8_t *d_mask, void *stream);

/* get_mask_by_color on an image that is HLS already (_utils.py:113-119 = cv2.inRange, inclusive
 * bounds, per channel): d_img is n frames of h x w 3-channel pixels (frame f at + f*frame_stride,
 * rows row_stride apart), d_mask n*h*w bytes, 255 where lo <= pixel <= hi.  lo/hi: 3 bytes, host. */
int melf_in_range(int device, const uint8_t *d_img, int n, int h, int w, size_t row_stride,
                  size_t frame_stride, const uint8_t *lo, const uint8_t *hi, uint8_t *d_mask, void *stream);

/* match_template (_utils.py:91-97): TM_CCOEFF of every frame's meter_rect crop against
 * the ctx template + minMaxLoc.  channels = 3: d_img is BGR and the lightness plane is
 * computed on the fly; channels = 1: d_img is an 8-bit single-channel image.
 * d_max_xy (n*2 int32), d_max_val (n float) and d_result (n * rh*rw float32, the
 * TM_CCOEFF map; may be NULL) are device buffers.  use_rect = 0 ignores meter_rect. */
int melf_match_template(melf_ctx *ctx, const uint8_t *d_img, int channels, int n, int h, int w,
                        size_t row_stride, size_t frame_stride, int use_rect,
                        int32_t *d_max_xy, float *d_max_val, float *d_result, void *stream);

/* melf_match_template for callers that cannot allocate: no counterpart in the reference.
 * Some callers cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it); this entry point calls no cudaMalloc / cudaFree and no allocating
 * library routine.  Same arguments, result and error codes as melf_match_template without d_result (arg-max
 * only).  The locate kernel's scratch (per-CTA lightness slabs + window sums, per-frame arg-max keys) lives in
 * a static device array of MELF_NOALLOC_SCRATCH_BYTES (about 40 MiB) that every GPU which loads the library
 * holds until exit.  Limits, each rejected with MELF_ERR_UNSUPPORTED and a message in melf_last_error:
 * n <= MELF_NOALLOC_MAX_FRAMES per call; a template the tensor-core kernel takes (<= 193 x 128 px).
 * Every call on a GPU SHARES that array: it is cleared on `stream` at the start of every call, and calls that
 * overlap on the GPU (different streams, or a captured graph launched twice concurrently) race -- serialise
 * them.  The array's device address is looked up once per device by melf_create (outside any capture). */
#define MELF_NOALLOC_MAX_FRAMES 8192
#define MELF_NOALLOC_SCRATCH_BYTES (148u * 282624u + MELF_NOALLOC_MAX_FRAMES * 8u)
int melf_match_template_noalloc(melf_ctx *ctx, const uint8_t *d_img, int channels, int n, int h, int w,
                                size_t row_stride, size_t frame_stride, int use_rect,
                                int32_t *d_max_xy, float *d_max_val, void *stream);

/* ---- ingest: cv2.imread of baseline JPEG streams on the GPU ------------------- */
/* ImageFile.get_bgr_image's cv2.imread (_image.py:49) for n JPEG streams held in HOST memory:
 * stream i is jpeg[offsets[i] .. offsets[i+1]).  The pixels are bit-identical to OpenCV's bundled
 * libjpeg-turbo (ISLOW IDCT, fancy upsampling).  Supported: 8-bit baseline/extended sequential
 * Huffman, one scan, YCbCr 4:4:4 / 4:2:2 / 4:2:0 or grey, with or without restart intervals (DRI);
 * 3-component streams that libjpeg reads as RGB (Adobe APP14 transform 0 or component ids 'R','G','B'
 * without a JFIF marker) are MELF_ERR_UNSUPPORTED; every stream of a
 * call must decode to h x w pixels with the same subsampling.  Anything else: MELF_ERR_UNSUPPORTED /
 * MELF_ERR_INVALID with a text in melf_last_error -- never a silently different image.
 * A stream whose entropy data ends early (file cut short) is decoded like libjpeg decodes it
 * (jdhuff.c insufficient_data): the MCU in which the data runs dry is finished on zero bits, the MCUs
 * after it stay mid-grey.  What is NOT restated is libjpeg's resynchronisation on the fake EOI marker
 * that its file source appends when such a stream also uses restart intervals; the Python layer
 * therefore sends every file without an EOI marker through cv2.imread.
 * d_bgr: device, frame f at d_bgr + f*frame_stride, rows row_stride apart.  Asynchronous on
 * `stream` once the streams have been copied into the library's pinned staging buffer. */
int melf_decode_jpeg(melf_ctx *ctx, const uint8_t *jpeg, const uint64_t *offsets, int n, int h, int w,
                     uint8_t *d_bgr, size_t row_stride, size_t frame_stride, void *stream);

/* Header probe of one stream (host only): size, component count, luma sampling factors.
 * Errors are reported through melf_last_error(NULL). */
int melf_jpeg_info(const uint8_t *jpeg, size_t len, int *h, int *w, int *components, int *h_samp,
                   int *v_samp);

/* Host-side batch helpers of get_meter_values' per-file preamble (_api.py:22-27, _image.py:46-51): file
 * sizes (-1 = cannot be opened), the files' bytes into one blob (ok[i] = 0: short read), and the header
 * probe of n streams at once: info[8i..] = {h, w, components, h_samp, v_samp, EXIF orientation, end-of-image
 * marker near the end (0/1), code} with code 0 = melf_read_jpeg_batch takes the stream, else the MELF_ERR_*
 * melf_jpeg_info returns (such files take the cv2.imread route).  No GPU involved; a few host threads. */
int melf_file_sizes(const char *const *paths, int n, int64_t *sizes);
int melf_read_files(const char *const *paths, int n, uint8_t *blob, const uint64_t *offsets, int32_t *ok);
int melf_jpeg_probe_batch(const uint8_t *jpeg, const uint64_t *offsets, int n, int32_t *info);

/* get_meter_values over JPEG files already read into memory (_api.py:22-33 without the file
 * I/O): decode, crop to meter_rect, locate and read, in pipelined chunks of up to 8192 streams on up
 * to 8 CUDA streams (MELF_JPEG_SLOTS / MELF_JPEG_CHUNK).
 * HOST streams in, HOST results out.  Synchronous. */
int melf_read_jpeg_batch(melf_ctx *ctx, const uint8_t *jpeg, const uint64_t *offsets, int n,
                         melf_result *out);

/* ---- offline calibration -------------------------------------------------- */
/* get_average_meter_image (_calibration.py:63-66): every frame translated so that its matched
 * dials rect (d_match_xy, n*2 int32 as written by melf_match_template) lands on
 * (anchor_x, anchor_y) = (30, 116) in the reference (_image.py:39-44, cv2.warpAffine with zero
 * border), then the running float64 mean of the /255 normalised frames in frame order
 * (_utils.py:64-88) and uint8(avg * 255 + 0.5).  d_out: h*w*3 bytes.  All buffers on `device`. */
int melf_average_translated(int device, const uint8_t *d_bgr, int n, int h, int w,
                            size_t row_stride, size_t frame_stride, const int32_t *d_match_xy,
                            int anchor_x, int anchor_y, uint8_t *d_out, void *stream);

/* Number of kernel launches issued by this library since load (bench bookkeeping). */
uint64_t melf_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* METERELF_B200_H */
