// C ABI of the dial-reading path (include/meterelf_b200.h): context, batching, launches.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "melf_common.cuh"

// ---- kernel launchers implemented in the other translation units ---------------------
#include "melf_args.h"

cudaError_t melf_launch_dials(const MelfDialsArgs &args, int n_frames, cudaStream_t stream);
dim3 melf_match_plan(MelfMatchArgs &a, int n_frames);
cudaError_t melf_launch_match(const MelfMatchArgs &args, dim3 grid, cudaStream_t stream);
cudaError_t melf_launch_match_unpack(const u64 *key, int n, int res_w, int32_t *xy, float *val,
                                     cudaStream_t stream);
#include "melf_jpeg.h"

cudaError_t melf_launch_convert(const MelfConvertArgs &a, int num_sms, cudaStream_t stream);
cudaError_t melf_launch_in_range(const uint8_t *img, size_t row_stride, size_t frame_stride, int n, int h, int w,
                                 uint32_t lo, uint32_t hi, uint8_t *mask, int num_sms, cudaStream_t stream);
bool melf_match_tc_supported(int tmpl_w, int tmpl_h);
size_t melf_match_tc_toeplitz_bytes(int tmpl_h);
void melf_match_tc_build_toeplitz(const uint8_t *tmpl, int tmpl_w, int tmpl_h, uint8_t *out);
long long melf_match_tc_plan(MelfMatchTcArgs &a, int n_frames);
size_t melf_match_tc_scratch_bytes(long long max_items, size_t *wsum_offset);
cudaError_t melf_launch_match_prep(const MelfMatchTcArgs &args, cudaStream_t stream);
cudaError_t melf_launch_match_tc(const MelfMatchTcArgs &args, int num_sms, cudaStream_t stream);
size_t melf_match_fused_scratch_bytes(int num_sms);
cudaError_t melf_launch_match_fused(const MelfMatchTcArgs &args, int num_sms, cudaStream_t stream);
size_t melf_match_fused_ov_scratch_bytes(int num_sms);
bool melf_match_small_supported(const MelfGeom &g, int tmpl_w, int tmpl_h, int channels);
void melf_match_small_plan(MelfSmallArgs &a);
size_t melf_match_small_tpos_bytes(const MelfSmallArgs &a);
size_t melf_match_small_wsum_bytes(const MelfSmallArgs &a);
void melf_match_small_build_tpos(const MelfSmallArgs &a, const uint8_t *tmpl, uint8_t *out);
cudaError_t melf_launch_match_small(const MelfSmallArgs &a, cudaStream_t stream);
cudaError_t melf_launch_match_fused_ov(const MelfMatchTcArgs &args, int num_sms, cudaStream_t stream);

namespace {

// Every entry point works on its own device and leaves the calling thread's current device as it
// found it (one process may drive several GPUs, and callers own buffers on "their" device).
struct DeviceScope {
    int prev = -1;
    cudaError_t status;
    explicit DeviceScope(int device)
    {
        status = cudaGetDevice(&prev);
        if (status == cudaSuccess && prev != device) status = cudaSetDevice(device); else if (status == cudaSuccess) prev = -1;
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceScope(const DeviceScope &) = delete;
    DeviceScope &operator=(const DeviceScope &) = delete;
};

std::atomic<uint64_t> g_launches{0};
std::mutex g_err_mutex;
std::string g_create_error;

constexpr int kChunkFrames = 512;      // frames per pipelined chunk in melf_read_batch
constexpr int kJpegChunkFrames = 8192; // most streams per pipelined chunk in melf_read_jpeg_batch ...
constexpr int kJpegSlots = 16;         // ... and (at most) chunks in flight: the Huffman stage runs one lane per stream
constexpr int kJpegSlotsDefault = 8;   // (MELF_JPEG_SLOTS overrides, up to kJpegSlots)
                                       // at low occupancy, so several chunks' kernels share the GPU
constexpr int kScratchSlots = 3 + kJpegSlots;   // [0] caller stream, [1], [2] melf_read_batch, then JPEG
constexpr long long kTcMaxItems = 4096; // work items per prep + correlation launch pair (~400 MB scratch)

}  // namespace

struct melf_ctx {
    int device = 0;
    int num_sms = 148;
    melf_params params{};              // pointers inside are NOT valid after create
    std::string err;

    // static device data
    MelfDialDev *d_dials = nullptr;
    uint16_t *d_ring_pos = nullptr;
    uint16_t *d_need_pos = nullptr;
    double *d_ring_angle = nullptr;
    double *d_ring_sqd = nullptr;
    uint32_t *d_tmpl_words = nullptr;
    uint32_t *d_ones_words = nullptr;
    uint8_t *d_toeplitz = nullptr;     // tensor-core path operands (NULL: template not supported there)
    bool use_tc = false;
    int pix_pitch = 0, pix_bytes = 0;       // per-dial shared-memory window of the dial kernel

    // melf_profile_*: CUDA events around every kernel launch while enabled
    bool profiling = false;
    struct Span { int kind; cudaEvent_t e0, e1; };
    std::vector<Span> spans;
    // small-result-map path (positions as M, frames as N): template per position for ONE crop geometry,
    // built on first use; frame operand + window sums per scratch slot
    std::vector<uint8_t> h_tmpl;
    const char *last_locate = "";    // which locate kernel the last call took (measurement: melf_last_locate)
    uint8_t *d_tpos = nullptr;
    int tpos_crop_w = -1, tpos_crop_h = -1;
    uint32_t *d_small_wsum[kScratchSlots] = {};
    size_t small_wsum_cap[kScratchSlots] = {};
    uint8_t *d_tc_scratch[kScratchSlots] = {};   // [0]: caller-stream path, [1], [2]: pipeline streams, ...
    long long tc_scratch_items[kScratchSlots] = {};
    long long sum_t = 0, n_t = 0;
    int tw4p = 0, steps = 0;
    std::vector<int32_t> win_x0, win_y0;

    // scratch for the async path
    u64 *d_keys = nullptr;
    int keys_cap = 0;
    MelfDialSlot *d_dial_slots[kScratchSlots] = {};   // per scratch slot, like d_tc_scratch
    int dial_slots_cap[kScratchSlots] = {};

    // pipeline resources of melf_read_batch
    cudaStream_t streams[2] = {nullptr, nullptr};
    uint8_t *d_chunk[2] = {nullptr, nullptr};
    size_t chunk_bytes = 0;
    melf_result *d_res[2] = {nullptr, nullptr};
    melf_result *h_res = nullptr;       // pinned, 2 * kChunkFrames
    u64 *d_chunk_keys[2] = {nullptr, nullptr};

    // JPEG ingest (melf_jpeg.cu): [0] caller-stream path, [1 ..] pipeline slots of melf_read_jpeg_batch
    MelfJpegState *jpeg[1 + kJpegSlots] = {};
    cudaStream_t jstreams[kJpegSlots] = {};
    cudaStream_t jcopy = nullptr;                 // the uploads of melf_read_jpeg_batch, in chunk order
    uint8_t *d_jchunk[kJpegSlots] = {};
    size_t jchunk_bytes = 0;
    melf_result *d_jres[kJpegSlots] = {};
    melf_result *h_jres = nullptr;      // pinned, kJpegSlots * kJpegChunkFrames
    u64 *d_jkeys[kJpegSlots] = {};
};

namespace {

int fail(melf_ctx *ctx, int code, const std::string &msg)
{
    if (ctx) ctx->err = msg;
    else {
        std::lock_guard<std::mutex> lock(g_err_mutex);
        g_create_error = msg;
    }
    return code;
}

int cuda_fail(melf_ctx *ctx, cudaError_t e, const char *what)
{
    return fail(ctx, MELF_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

#define MELF_CUDA(ctx, call)                                         \
    do {                                                             \
        cudaError_t e_ = (call);                                     \
        if (e_ != cudaSuccess) return cuda_fail((ctx), e_, #call);   \
    } while (0)

// runs `launch` (a kernel launch on `stream`), bracketed by events of kind `kind` while profiling
template <typename F>
cudaError_t timed(melf_ctx *ctx, int kind, cudaStream_t stream, F launch)
{
    if (!ctx->profiling) return launch();
    melf_ctx::Span sp{kind, nullptr, nullptr};
    cudaError_t e = cudaEventCreate(&sp.e0);
    if (e != cudaSuccess) return e;
    e = cudaEventCreate(&sp.e1);
    if (e != cudaSuccess) return e;
    cudaEventRecord(sp.e0, stream);
    e = launch();
    cudaEventRecord(sp.e1, stream);
    ctx->spans.push_back(sp);
    return e;
}

// numpy-slice clipping of the meter_rect against the frame (_utils.py:52-54)
int make_geom(melf_ctx *ctx, int h, int w, bool use_rect, MelfGeom *g)
{
    const melf_params &P = ctx->params;
    int x0 = 0, y0 = 0, x1 = w, y1 = h;
    if (use_rect) {
        x0 = P.rect_x0 < w ? P.rect_x0 : w; x1 = P.rect_x1 < w ? P.rect_x1 : w;
        y0 = P.rect_y0 < h ? P.rect_y0 : h; y1 = P.rect_y1 < h ? P.rect_y1 : h;
    }
    g->crop_x0 = x0; g->crop_y0 = y0;
    g->crop_w = x1 - x0; g->crop_h = y1 - y0;
    g->res_w = g->crop_w - P.tmpl_w + 1;
    g->res_h = g->crop_h - P.tmpl_h + 1;
    if (g->res_w < 1 || g->res_h < 1)
        return fail(ctx, MELF_ERR_TOO_SMALL, "meter_rect crop is smaller than the dials template");
    return MELF_OK;
}

void fill_match_args(const melf_ctx *ctx, MelfMatchArgs *a)
{
    a->tmpl_words = ctx->d_tmpl_words;
    a->ones_words = ctx->d_ones_words;
    a->tmpl_w = ctx->params.tmpl_w;
    a->tmpl_h = ctx->params.tmpl_h;
    a->sum_t = ctx->sum_t;
    a->n_t = ctx->n_t;
}

// TM_CCOEFF arg-max keys for n frames: tensor-core kernel when the template fits it, else
// the dp4a kernel (both CUDA, both exact; MELF_MATCH_IMPL=dp4a forces the latter)
int launch_locate(melf_ctx *ctx, const uint8_t *d_img, int channels, int n, int h, int w,
                  size_t row_stride, size_t frame_stride, const MelfGeom &geom, u64 *d_keys,
                  float *d_result, cudaStream_t stream, int scratch_slot)
{
    MELF_CUDA(ctx, cudaMemsetAsync(d_keys, 0, (size_t)n * sizeof(u64), stream));
    // small result maps (sample set 2): one GEMM per group of 32 frames, positions as M (melf_match_tc.cu, last
    // section).  Its time barely depends on the batch (one wave of 32-frame groups), the Toeplitz kernel's is
    // proportional to it, so a cost model in clocks picks the path: measured per MMA slot 132 clk in the fused
    // GEMM kernel (+ 35 clk per frame for the edge kernel) against ~78 clk per MMA in the Toeplitz kernel
    // (profiles/r02_small_launches.csv).  MELF_MATCH_SMALL=0 / 1 forces the choice.
    bool small_path = false;
    if (ctx->use_tc && !d_result && melf_match_small_supported(geom, ctx->params.tmpl_w, ctx->params.tmpl_h, channels)) {
        static const int forced = std::getenv("MELF_MATCH_SMALL") ? std::atoi(std::getenv("MELF_MATCH_SMALL")) : -1;
        if (forced >= 0) {
            small_path = forced != 0;
        } else {
            MelfSmallArgs plan{};
            plan.geom = geom; plan.tmpl_w = ctx->params.tmpl_w; plan.tmpl_h = ctx->params.tmpl_h;
            melf_match_small_plan(plan);
            const long long groups = ((long long)n + 31) / 32, sms = ctx->num_sms;
            const long long t_small = ((groups + sms - 1) / sms) * (long long)plan.n_kstages * plan.n_mt * 2 * 132 + 35ll * n;
            const long long n_pairs = (ctx->params.tmpl_h + 1) / 2;
            const long long kpp = std::min(8, (ctx->params.tmpl_w + 63 + 31) / 32);      // K-steps per pair that are not all zero
            const long long t_toeplitz = (((long long)n + 2 * sms - 1) / (2 * sms)) * (n_pairs * 2 * kpp * 78 + 4000);
            small_path = t_small < t_toeplitz;
        }
    }
    if (small_path) {
        ctx->last_locate = "melf_small_fused_kernel";
        MelfSmallArgs a{};
        a.img = d_img;
        const uint8_t *end = d_img + (size_t)(n - 1) * frame_stride + (size_t)(h - 1) * row_stride + (size_t)w * channels;
        a.img_limit = reinterpret_cast<const uint8_t *>((reinterpret_cast<uintptr_t>(end) + 15) & ~(uintptr_t)15);
        a.row_stride = row_stride; a.frame_stride = frame_stride; a.channels = channels; a.geom = geom;
        a.n_frames = n; a.tmpl_w = ctx->params.tmpl_w; a.tmpl_h = ctx->params.tmpl_h;
        a.sum_t = ctx->sum_t; a.n_t = ctx->n_t; a.match_key = d_keys;
        melf_match_small_plan(a);
        if (ctx->tpos_crop_w != geom.crop_w || ctx->tpos_crop_h != geom.crop_h) {
            // (kernels of earlier calls may still read the old image)
            MELF_CUDA(ctx, cudaDeviceSynchronize());
            if (ctx->d_tpos) cudaFree(ctx->d_tpos);
            ctx->d_tpos = nullptr; ctx->tpos_crop_w = ctx->tpos_crop_h = -1;
            std::vector<uint8_t> tpos(melf_match_small_tpos_bytes(a));
            melf_match_small_build_tpos(a, ctx->h_tmpl.data(), tpos.data());
            MELF_CUDA(ctx, cudaMalloc(&ctx->d_tpos, tpos.size()));
            MELF_CUDA(ctx, cudaMemcpy(ctx->d_tpos, tpos.data(), tpos.size(), cudaMemcpyHostToDevice));
            ctx->tpos_crop_w = geom.crop_w; ctx->tpos_crop_h = geom.crop_h;
        }
        a.tpos = ctx->d_tpos;
        const size_t wsum_need = melf_match_small_wsum_bytes(a);
        if (ctx->small_wsum_cap[scratch_slot] < wsum_need) {
            if (ctx->d_small_wsum[scratch_slot]) { MELF_CUDA(ctx, cudaStreamSynchronize(stream)); cudaFree(ctx->d_small_wsum[scratch_slot]); }
            ctx->d_small_wsum[scratch_slot] = nullptr; ctx->small_wsum_cap[scratch_slot] = 0;
            MELF_CUDA(ctx, cudaMalloc(reinterpret_cast<void **>(&ctx->d_small_wsum[scratch_slot]), wsum_need));
            ctx->small_wsum_cap[scratch_slot] = wsum_need;
        }
        a.wsum = ctx->d_small_wsum[scratch_slot];
        MELF_CUDA(ctx, timed(ctx, MELF_PROF_MATCH_TC, stream, [&] { return melf_launch_match_small(a, stream); }));
        g_launches += 2;
        return MELF_OK;
    }
    if (ctx->use_tc) {
        ctx->last_locate = "melf_match_fused_ov_kernel";
        MelfMatchTcArgs t{};
        t.img = d_img;
        const uint8_t *end = d_img + (size_t)(n - 1) * frame_stride + (size_t)(h - 1) * row_stride +
                             (size_t)w * channels;
        t.img_limit = reinterpret_cast<const uint8_t *>((reinterpret_cast<uintptr_t>(end) + 15) & ~(uintptr_t)15);
        t.row_stride = row_stride; t.frame_stride = frame_stride;
        t.channels = channels;
        t.img_w = w; t.img_h = h; t.img_n = n;
        t.geom = geom;
        t.toeplitz = ctx->d_toeplitz;
        t.tmpl_w = ctx->params.tmpl_w; t.tmpl_h = ctx->params.tmpl_h;
        t.sum_t = ctx->sum_t; t.n_t = ctx->n_t;
        t.match_key = d_keys;
        t.result = d_result;
        melf_match_tc_plan(t, n);
        // MELF_MATCH_TWO_KERNEL=1 selects the older preparation + correlation kernel pair (A/B runs)
        static const bool two_kernel = std::getenv("MELF_MATCH_TWO_KERNEL") && std::atoi(std::getenv("MELF_MATCH_TWO_KERNEL"));
        if (two_kernel) {
            long long cap = t.n_items < kTcMaxItems ? t.n_items : kTcMaxItems;
            cap += cap & 1;
            if (ctx->tc_scratch_items[scratch_slot] < cap) {
                if (ctx->d_tc_scratch[scratch_slot]) cudaFree(ctx->d_tc_scratch[scratch_slot]);
                ctx->d_tc_scratch[scratch_slot] = nullptr; ctx->tc_scratch_items[scratch_slot] = 0;
                MELF_CUDA(ctx, cudaMalloc(&ctx->d_tc_scratch[scratch_slot], melf_match_tc_scratch_bytes(cap, nullptr)));
                ctx->tc_scratch_items[scratch_slot] = cap;
            }
            size_t wsum_off = 0;
            melf_match_tc_scratch_bytes(ctx->tc_scratch_items[scratch_slot], &wsum_off);
            t.slabs = ctx->d_tc_scratch[scratch_slot];
            t.wsum = reinterpret_cast<uint32_t *>(ctx->d_tc_scratch[scratch_slot] + wsum_off);
            t.max_items = ctx->tc_scratch_items[scratch_slot];
        } else {
            // fused kernel: per-CTA scratch (slabs of the next pass + window sums of the current one)
            if (ctx->tc_scratch_items[scratch_slot] != -1) {
                if (ctx->d_tc_scratch[scratch_slot]) cudaFree(ctx->d_tc_scratch[scratch_slot]);
                ctx->d_tc_scratch[scratch_slot] = nullptr; ctx->tc_scratch_items[scratch_slot] = 0;
                const size_t bytes = melf_match_fused_ov_scratch_bytes(ctx->num_sms);     // the larger of the two layouts
                MELF_CUDA(ctx, cudaMalloc(&ctx->d_tc_scratch[scratch_slot], bytes));
                // on the stream the kernels run on (a memset on the legacy stream is not ordered against
                // a cudaStreamNonBlocking stream)
                MELF_CUDA(ctx, cudaMemsetAsync(ctx->d_tc_scratch[scratch_slot], 0, bytes, stream));
                ctx->tc_scratch_items[scratch_slot] = -1;         // marks "fused layout"
            }
            t.slabs = ctx->d_tc_scratch[scratch_slot];
            t.max_items = t.n_items;
        }
        // MELF_TC_TIMELINE=<file>: debug dump of CTA 0's per-pass clock64 stamps (synchronises)
        const char *tl_path = std::getenv("MELF_TC_TIMELINE");
        long long *d_tl = nullptr;
        if (tl_path) {
            MELF_CUDA(ctx, cudaMalloc(&d_tl, 64 * sizeof(long long)));
            MELF_CUDA(ctx, cudaMemsetAsync(d_tl, 0, 64 * sizeof(long long), stream));
            t.timeline = d_tl;
        }
        if (const char *skip = std::getenv("MELF_TC_DEBUG_SKIP")) t.debug_skip = std::atoi(skip);
        // default: drain of pass k under the MMAs of pass k + 1 (three TMEM accumulators);
        // MELF_MATCH_OVERLAP=0 selects the kernel with the drain between the passes (A/B runs)
        static const bool overlap = !(std::getenv("MELF_MATCH_OVERLAP") && !std::atoi(std::getenv("MELF_MATCH_OVERLAP")));
        if (!two_kernel) {
            MELF_CUDA(ctx, timed(ctx, MELF_PROF_MATCH_TC, stream, [&] {
                return overlap ? melf_launch_match_fused_ov(t, ctx->num_sms, stream)
                               : melf_launch_match_fused(t, ctx->num_sms, stream); }));
            g_launches += 1;
        } else
        for (t.item0 = 0; t.item0 < t.n_items; t.item0 += t.max_items) {
            MELF_CUDA(ctx, timed(ctx, MELF_PROF_MATCH_PREP, stream, [&] { return melf_launch_match_prep(t, stream); }));
            MELF_CUDA(ctx, timed(ctx, MELF_PROF_MATCH_TC, stream, [&] { return melf_launch_match_tc(t, ctx->num_sms, stream); }));
            g_launches += 2;
        }
        if (d_tl) {
            long long h_tl[64];
            MELF_CUDA(ctx, cudaStreamSynchronize(stream));
            MELF_CUDA(ctx, cudaMemcpy(h_tl, d_tl, sizeof h_tl, cudaMemcpyDeviceToHost));
            cudaFree(d_tl);
            if (FILE *fp = std::fopen(tl_path, "w")) {
                for (int i = 0; i < 8; i++) {
                    for (int j = 0; j < 8; j++) std::fprintf(fp, "%lld ", h_tl[8 * i + j] - ((j < 7 || !two_kernel) ? h_tl[0] : 0));
                    std::fprintf(fp, "\n");
                }
                std::fclose(fp);
            }
        }
        return MELF_OK;
    }
    ctx->last_locate = "melf_match_kernel";
    MelfMatchArgs m{};
    fill_match_args(ctx, &m);
    m.channels = channels;
    m.row_stride = row_stride; m.frame_stride = frame_stride;
    m.geom = geom;
    // grid.z is limited to 65535 frames per launch
    for (int f0 = 0; f0 < n; f0 += 65535) {
        const int nf = n - f0 < 65535 ? n - f0 : 65535;
        MelfMatchArgs mm = m;
        mm.img = d_img + (size_t)f0 * frame_stride;
        mm.match_key = d_keys + f0;
        mm.result = d_result ? d_result + (size_t)f0 * geom.res_w * geom.res_h : nullptr;
        const dim3 grid = melf_match_plan(mm, nf);
        MELF_CUDA(ctx, timed(ctx, MELF_PROF_MATCH_DP4A, stream, [&] { return melf_launch_match(mm, grid, stream); }));
        g_launches++;
    }
    return MELF_OK;
}

int ensure_keys(melf_ctx *ctx, int n)
{
    if (n <= ctx->keys_cap) return MELF_OK;
    if (ctx->d_keys) cudaFree(ctx->d_keys);
    ctx->d_keys = nullptr; ctx->keys_cap = 0;
    MELF_CUDA(ctx, cudaMalloc(&ctx->d_keys, (size_t)n * sizeof(u64)));
    ctx->keys_cap = n;
    return MELF_OK;
}

// locate + read on device-resident frames; keys = per-frame scratch (n entries)
int run_path(melf_ctx *ctx, const uint8_t *d_bgr, int n, int frame_h, int frame_w,
             size_t row_stride, size_t frame_stride, const MelfGeom &geom, u64 *d_keys, melf_result *d_out, const melf_taps *taps,
             cudaStream_t stream, int scratch_slot)
{
    if (ctx->params.n_dials < 1) return fail(ctx, MELF_ERR_INVALID, "locate-only context (no dials): nothing to read");
    int rc = launch_locate(ctx, d_bgr, 3, n, frame_h, frame_w, row_stride, frame_stride, geom, d_keys,
                           nullptr, stream, scratch_slot);
    if (rc) return rc;
    if (ctx->dial_slots_cap[scratch_slot] < n) {
        if (ctx->d_dial_slots[scratch_slot]) cudaFree(ctx->d_dial_slots[scratch_slot]);
        ctx->d_dial_slots[scratch_slot] = nullptr; ctx->dial_slots_cap[scratch_slot] = 0;
        MELF_CUDA(ctx, cudaMalloc(&ctx->d_dial_slots[scratch_slot], (size_t)n * sizeof(MelfDialSlot)));
        MELF_CUDA(ctx, cudaMemsetAsync(ctx->d_dial_slots[scratch_slot], 0, (size_t)n * sizeof(MelfDialSlot), stream));
        ctx->dial_slots_cap[scratch_slot] = n;
    }
    MelfDialsArgs a{};
    a.slots = ctx->d_dial_slots[scratch_slot]; a.n_frames = n;
    a.bgr = d_bgr; a.row_stride = row_stride; a.frame_stride = frame_stride;
    a.bgr_end = d_bgr + (size_t)(n - 1) * frame_stride + (size_t)(frame_h - 1) * row_stride + (size_t)frame_w * 3;
    a.geom = geom;
    a.dials = ctx->d_dials;
    a.need_pos = ctx->d_need_pos; a.ring_pos = ctx->d_ring_pos; a.ring_angle = ctx->d_ring_angle; a.ring_sqd = ctx->d_ring_sqd;
    a.match_key = d_keys;
    a.pix_pitch = ctx->pix_pitch; a.pix_bytes = ctx->pix_bytes;
    a.out = d_out;
    if (taps) a.taps = *taps; else std::memset(&a.taps, 0, sizeof a.taps);
    a.threshold = ctx->params.match_threshold;
    a.n_dials = ctx->params.n_dials;
    a.tmpl_w = ctx->params.tmpl_w; a.tmpl_h = ctx->params.tmpl_h;
    a.hue_shift = ctx->params.hue_shift; a.s_tail = ctx->params.s_tail;
    for (int k = 0; k < 4; k++) a.value_order[k] = ctx->params.value_order[k];
    // MELF_DIALS_TIMELINE=<file>: debug dump of per-phase clock64 stamps (synchronises)
    const char *tl_path = std::getenv("MELF_DIALS_TIMELINE");
    long long *d_tl = nullptr;
    if (tl_path) {
        MELF_CUDA(ctx, cudaMalloc(&d_tl, 64 * sizeof(long long)));
        MELF_CUDA(ctx, cudaMemsetAsync(d_tl, 0, 64 * sizeof(long long), stream));
        a.timeline = d_tl;
    }
    MELF_CUDA(ctx, timed(ctx, MELF_PROF_DIALS, stream, [&] { return melf_launch_dials(a, n, stream); }));
    g_launches++;
    if (d_tl) {
        long long h_tl[64];
        MELF_CUDA(ctx, cudaStreamSynchronize(stream));
        MELF_CUDA(ctx, cudaMemcpy(h_tl, d_tl, sizeof h_tl, cudaMemcpyDeviceToHost));
        cudaFree(d_tl);
        if (FILE *fp = std::fopen(tl_path, "w")) {
            for (int i = 0; i < 8; i++) {
                for (int j = 0; j < 7; j++) std::fprintf(fp, "%lld ", h_tl[8 * i + j] - h_tl[8 * i]);
                std::fprintf(fp, "%lld\n", h_tl[8 * i + 7]);     // 32-pixel groups converted in full
            }
            std::fclose(fp);
        }
    }
    return MELF_OK;
}


// event pairs for the five stages of one JPEG decode while profiling (NULL otherwise)
struct JpegSpans {
    cudaEvent_t ev[5][2];
    bool on = false;
    explicit JpegSpans(melf_ctx *ctx)
    {
        if (!ctx->profiling) return;
        on = true;
        for (auto &p : ev) { cudaEventCreate(&p[0]); cudaEventCreate(&p[1]); }
    }
    cudaEvent_t (*get())[2] { return on ? ev : nullptr; }
    void commit(melf_ctx *ctx)
    {
        if (!on) return;
        for (int k = 0; k < 5; k++) ctx->spans.push_back(melf_ctx::Span{MELF_PROF_JPEG_UPLOAD + k, ev[k][0], ev[k][1]});
    }
};

void free_ctx(melf_ctx *ctx)
{
    if (!ctx) return;
    DeviceScope dev_scope(ctx->device);
    cudaFree(ctx->d_dials); cudaFree(ctx->d_ring_pos); cudaFree(ctx->d_need_pos); cudaFree(ctx->d_ring_angle);
    cudaFree(ctx->d_ring_sqd); cudaFree(ctx->d_tmpl_words); cudaFree(ctx->d_ones_words);
    cudaFree(ctx->d_toeplitz);
    cudaFree(ctx->d_tpos);
    for (int i = 0; i < kScratchSlots; i++) cudaFree(ctx->d_small_wsum[i]);
    cudaFree(ctx->d_keys);
    for (int i = 0; i < kScratchSlots; i++) cudaFree(ctx->d_dial_slots[i]);
    for (int i = 0; i < kScratchSlots; i++) cudaFree(ctx->d_tc_scratch[i]);
    for (int i = 0; i < 2; i++) {
        if (ctx->streams[i]) cudaStreamDestroy(ctx->streams[i]);
        cudaFree(ctx->d_chunk[i]); cudaFree(ctx->d_res[i]); cudaFree(ctx->d_chunk_keys[i]);
    }
    if (ctx->h_res) cudaFreeHost(ctx->h_res);
    for (int i = 0; i < 1 + kJpegSlots; i++) melf_jpeg_state_free(ctx->jpeg[i]);
    for (int i = 0; i < kJpegSlots; i++) {
        if (ctx->jstreams[i]) cudaStreamDestroy(ctx->jstreams[i]);
        cudaFree(ctx->d_jchunk[i]); cudaFree(ctx->d_jres[i]); cudaFree(ctx->d_jkeys[i]);
    }
    if (ctx->h_jres) cudaFreeHost(ctx->h_jres);
    if (ctx->jcopy) cudaStreamDestroy(ctx->jcopy);
    delete ctx;
}

}  // namespace

// ---- melf_match_template_noalloc: scratch that is never allocated ----------------------------------------
// The locate kernel's scratch for callers that cannot call cudaMalloc (CUDA graph capture, serving loops):
// [per-CTA slabs + window sums of melf_match_fused_ov_kernel for kNoAllocCtas CTAs][kNoAllocMaxFrames arg-max keys].
// One copy per GPU that loads the library, held until exit.  In HOST code the array's name is a host-side
// stand-in, not a device address: noalloc_setup() asks the runtime for the address, once per device.
constexpr int kNoAllocCtas = 148;                                   // B200: one CTA per SM
constexpr size_t kNoAllocCtaBytes = 282624;                         // F_SCRATCH_PER_CTA_OV (checked at run time)
constexpr size_t kNoAllocKeysOff = (size_t)kNoAllocCtas * kNoAllocCtaBytes;              // 41,828,352
constexpr size_t kNoAllocBytes = kNoAllocKeysOff + (size_t)MELF_NOALLOC_MAX_FRAMES * sizeof(u64);
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

namespace {

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mutex;
unsigned char *g_noalloc_ptr[kNoAllocMaxDevices] = {};               // device address of g_noalloc_scratch per device
cudaError_t g_noalloc_err[kNoAllocMaxDevices] = {};

// Once per device, with that device current (the caller holds a DeviceScope): the device address of the array.
cudaError_t noalloc_setup(int device, unsigned char **ptr)
{
    if (device < 0 || device >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device] && g_noalloc_err[device] == cudaSuccess) {
        void *p = nullptr;
        g_noalloc_err[device] = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (g_noalloc_err[device] == cudaSuccess) g_noalloc_ptr[device] = static_cast<unsigned char *>(p);
    }
    if (ptr) *ptr = g_noalloc_ptr[device];
    return g_noalloc_err[device];
}

}  // namespace

extern "C" {

int melf_abi_version(void) { return MELF_ABI_VERSION; }

uint64_t melf_launch_count(void) { return g_launches.load(); }

const char *melf_last_error(const melf_ctx *ctx)
{
    if (ctx) return ctx->err.c_str();
    static thread_local std::string copy;
    std::lock_guard<std::mutex> lock(g_err_mutex);
    copy = g_create_error;
    return copy.c_str();
}

int melf_create(melf_ctx **out, int device, const melf_params *P)
{
    if (!out) return fail(nullptr, MELF_ERR_INVALID, "out is NULL");
    *out = nullptr;
    std::string err;
    int rc = melf_host_check_params(P, &err);
    if (rc) return fail(nullptr, rc, err);

    // host-side static tables first (no CUDA needed to fail on bad params)
    std::vector<MelfDialDev> dials(P->n_dials);
    MelfRingTables rings;
    for (int d = 0; d < P->n_dials; d++) {
        rc = melf_host_build_dial(P, d, &dials[d], &rings, &err);
        if (rc) return fail(nullptr, rc, err);
    }

    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceCount");
    if (device < 0 || device >= count)
        return fail(nullptr, MELF_ERR_CUDA, "no such CUDA device");
    DeviceScope dev_scope(device);
    if (dev_scope.status != cudaSuccess) return cuda_fail(nullptr, dev_scope.status, "cudaSetDevice");

    melf_ctx *ctx = new melf_ctx();
    ctx->device = device;
    ctx->params = *P;
    ctx->h_tmpl.assign(P->tmpl, P->tmpl + (size_t)P->tmpl_w * P->tmpl_h);
    ctx->params.tmpl = nullptr;
    for (int d = 0; d < MELF_MAX_DIALS; d++) {
        ctx->params.dials[d].mask = nullptr;
        ctx->params.dials[d].circle_mask = nullptr;
    }
    cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, device);
    {
        int max_w = 1, max_h = 1, max_ring = 0;
        for (int d = 0; d < P->n_dials; d++) {
            ctx->win_x0.push_back(dials[d].win_x0);
            ctx->win_y0.push_back(dials[d].win_y0);
            max_w = dials[d].win_w > max_w ? dials[d].win_w : max_w;
            max_h = dials[d].win_h > max_h ? dials[d].win_h : max_h;
            max_ring = dials[d].n_ring > max_ring ? dials[d].n_ring : max_ring;
        }
        ctx->pix_pitch = ((3 * max_w + 15 + 15) / 16) * 16;          // 3 w bytes at any 16-byte phase
        const int lists = 512 + ((2 * max_ring + 15) / 16) * 16;     // momentum buffer + kept-index list
        ctx->pix_bytes = max_h * ctx->pix_pitch > lists ? max_h * ctx->pix_pitch : lists;
    }

    // template, packed 4 columns per word in the layout the match kernel streams:
    // words[(jp*steps + s)*2 + q] = columns 4*(2jp+q)..+3 of template row s (0 beyond)
    {
        MelfMatchArgs plan{};
        plan.tmpl_w = P->tmpl_w; plan.tmpl_h = P->tmpl_h;
        plan.geom.res_w = 1; plan.geom.res_h = 1;
        melf_match_plan(plan, 1);
        ctx->tw4p = plan.tw4p; ctx->steps = plan.steps;
        std::vector<uint32_t> words((size_t)plan.tw4p * plan.steps, 0), ones(plan.tw4p, 0);
        for (int j4 = 0; j4 < plan.tw4p; j4++) {
            for (int b = 0; b < 4; b++)
                if (4 * j4 + b < P->tmpl_w) ones[j4] |= 1u << (8 * b);
            for (int s = 0; s < plan.steps; s++) {
                uint32_t w = 0;
                if (s < P->tmpl_h)
                    for (int b = 0; b < 4; b++)
                        if (4 * j4 + b < P->tmpl_w)
                            w |= (uint32_t)P->tmpl[(size_t)s * P->tmpl_w + 4 * j4 + b] << (8 * b);
                words[((size_t)(j4 >> 1) * plan.steps + s) * 2 + (j4 & 1)] = w;
            }
        }
        ctx->n_t = (long long)P->tmpl_w * P->tmpl_h;
        ctx->sum_t = 0;
        for (long long i = 0; i < ctx->n_t; i++) ctx->sum_t += P->tmpl[i];

#define MELF_CREATE_CUDA(call)                                                  \
    do {                                                                        \
        cudaError_t e_ = (call);                                                \
        if (e_ != cudaSuccess) {                                                \
            int rc_ = cuda_fail(nullptr, e_, #call);                            \
            free_ctx(ctx);                                                      \
            return rc_;                                                         \
        }                                                                       \
    } while (0)

        MELF_CREATE_CUDA(cudaMalloc(&ctx->d_tmpl_words, words.size() * 4));
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_tmpl_words, words.data(), words.size() * 4, cudaMemcpyHostToDevice));
        MELF_CREATE_CUDA(cudaMalloc(&ctx->d_ones_words, ones.size() * 4));
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_ones_words, ones.data(), ones.size() * 4, cudaMemcpyHostToDevice));
    }
    noalloc_setup(device, nullptr);       // address of the static scratch of melf_match_template_noalloc, once per device
    if (melf_match_tc_supported(P->tmpl_w, P->tmpl_h)) {
        std::vector<uint8_t> toep(melf_match_tc_toeplitz_bytes(P->tmpl_h));
        melf_match_tc_build_toeplitz(P->tmpl, P->tmpl_w, P->tmpl_h, toep.data());
        MELF_CREATE_CUDA(cudaMalloc(&ctx->d_toeplitz, toep.size()));
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_toeplitz, toep.data(), toep.size(), cudaMemcpyHostToDevice));
        const char *impl = std::getenv("MELF_MATCH_IMPL");
        ctx->use_tc = !(impl && std::strcmp(impl, "dp4a") == 0);
    }
    MELF_CREATE_CUDA(cudaMalloc(&ctx->d_dials, (dials.size() ? dials.size() : 1) * sizeof(MelfDialDev)));
    if (!dials.empty())
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_dials, dials.data(), dials.size() * sizeof(MelfDialDev), cudaMemcpyHostToDevice));
    {
        const size_t nn = rings.need.size() ? rings.need.size() : 1;
        MELF_CREATE_CUDA(cudaMalloc(&ctx->d_need_pos, nn * sizeof(uint16_t)));
        if (!rings.need.empty())
            MELF_CREATE_CUDA(cudaMemcpy(ctx->d_need_pos, rings.need.data(), nn * sizeof(uint16_t), cudaMemcpyHostToDevice));
    }
    const size_t nr = rings.pos.size() ? rings.pos.size() : 1;
    MELF_CREATE_CUDA(cudaMalloc(&ctx->d_ring_pos, nr * sizeof(uint16_t)));
    MELF_CREATE_CUDA(cudaMalloc(&ctx->d_ring_angle, nr * sizeof(double)));
    MELF_CREATE_CUDA(cudaMalloc(&ctx->d_ring_sqd, nr * sizeof(double)));
    if (!rings.pos.empty()) {
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_ring_pos, rings.pos.data(), nr * sizeof(uint16_t), cudaMemcpyHostToDevice));
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_ring_angle, rings.angle.data(), nr * sizeof(double), cudaMemcpyHostToDevice));
        MELF_CREATE_CUDA(cudaMemcpy(ctx->d_ring_sqd, rings.sqd.data(), nr * sizeof(double), cudaMemcpyHostToDevice));
    }
    *out = ctx;
    return MELF_OK;
}

void melf_destroy(melf_ctx *ctx) { free_ctx(ctx); }

const char *melf_last_locate(const melf_ctx *ctx) { return ctx ? ctx->last_locate : ""; }

int melf_profile_enable(melf_ctx *ctx, int on)
{
    if (!ctx) return MELF_ERR_INVALID;
    ctx->profiling = on != 0;
    return MELF_OK;
}

int melf_profile_read(melf_ctx *ctx, float *ms, int32_t *launches)
{
    if (!ctx || !ms || !launches) return MELF_ERR_INVALID;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    MELF_CUDA(ctx, cudaDeviceSynchronize());
    for (int k = 0; k < MELF_PROF_KINDS; k++) { ms[k] = 0.f; launches[k] = 0; }
    // MELF_PROF_TRACE=<file>: every span as "kind start_ms end_ms" relative to the first one (debug timeline)
    if (const char *trace = std::getenv("MELF_PROF_TRACE")) {
        if (FILE *fp = ctx->spans.empty() ? nullptr : std::fopen(trace, "w")) {
            for (const melf_ctx::Span &sp : ctx->spans) {
                float t0 = 0.f, t1 = 0.f;
                cudaEventElapsedTime(&t0, ctx->spans.front().e0, sp.e0);
                cudaEventElapsedTime(&t1, ctx->spans.front().e0, sp.e1);
                std::fprintf(fp, "%d %.4f %.4f\n", sp.kind, t0, t1);
            }
            std::fclose(fp);
        }
    }
    for (const melf_ctx::Span &sp : ctx->spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, sp.e0, sp.e1) == cudaSuccess && sp.kind >= 0 && sp.kind < MELF_PROF_KINDS) {
            ms[sp.kind] += t;
            launches[sp.kind]++;
        }
        cudaEventDestroy(sp.e0);
        cudaEventDestroy(sp.e1);
    }
    ctx->spans.clear();
    return MELF_OK;
}

int melf_dial_window(const melf_ctx *ctx, int dial, int32_t *win_x0, int32_t *win_y0)
{
    if (!ctx || dial < 0 || dial >= ctx->params.n_dials || !win_x0 || !win_y0) return MELF_ERR_INVALID;
    *win_x0 = ctx->win_x0[dial];
    *win_y0 = ctx->win_y0[dial];
    return MELF_OK;
}

int melf_read_batch_async(melf_ctx *ctx, const uint8_t *d_bgr, int n, int h, int w,
                          size_t row_stride, size_t frame_stride,
                          melf_result *d_out, const melf_taps *taps, void *stream)
{
    if (!ctx) return MELF_ERR_INVALID;
    if (n < 0 || h < 1 || w < 1 || !d_out || (n > 0 && !d_bgr))
        return fail(ctx, MELF_ERR_INVALID, "bad batch arguments");
    if (row_stride < (size_t)w * 3) return fail(ctx, MELF_ERR_INVALID, "row_stride < 3*w");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    MelfGeom geom;
    int rc = make_geom(ctx, h, w, true, &geom);
    if (rc) return rc;
    rc = ensure_keys(ctx, n);
    if (rc) return rc;
    return run_path(ctx, d_bgr, n, h, w, row_stride, frame_stride, geom, ctx->d_keys, d_out, taps,
                    (cudaStream_t)stream, 0);
}

int melf_read_batch(melf_ctx *ctx, const uint8_t *bgr, int n, int h, int w,
                    size_t row_stride, size_t frame_stride, melf_result *out)
{
    if (!ctx) return MELF_ERR_INVALID;
    if (n < 0 || h < 1 || w < 1 || (n > 0 && (!bgr || !out)))
        return fail(ctx, MELF_ERR_INVALID, "bad batch arguments");
    if (row_stride < (size_t)w * 3) return fail(ctx, MELF_ERR_INVALID, "row_stride < 3*w");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    MelfGeom geom;
    int rc = make_geom(ctx, h, w, true, &geom);
    if (rc) return rc;

    // only the meter_rect rows travel: compact device layout [frame][crop_h][pitch]
    const size_t crop_row_bytes = (size_t)geom.crop_w * 3;
    const size_t pitch = (crop_row_bytes + 15) & ~(size_t)15;
    const size_t frame_bytes = pitch * geom.crop_h;
    const size_t need = frame_bytes * kChunkFrames;
    if (!ctx->streams[0]) {
        for (int i = 0; i < 2; i++) {
            MELF_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->streams[i], cudaStreamNonBlocking));
            MELF_CUDA(ctx, cudaMalloc(&ctx->d_res[i], sizeof(melf_result) * kChunkFrames));
            MELF_CUDA(ctx, cudaMalloc(&ctx->d_chunk_keys[i], sizeof(u64) * kChunkFrames));
        }
        MELF_CUDA(ctx, cudaMallocHost(&ctx->h_res, sizeof(melf_result) * kChunkFrames * 2));
    }
    if (need > ctx->chunk_bytes) {
        for (int i = 0; i < 2; i++) {
            if (ctx->d_chunk[i]) cudaFree(ctx->d_chunk[i]);
            ctx->d_chunk[i] = nullptr;
        }
        ctx->chunk_bytes = 0;
        for (int i = 0; i < 2; i++) MELF_CUDA(ctx, cudaMalloc(&ctx->d_chunk[i], need));
        ctx->chunk_bytes = need;
    }
    MelfGeom cg = geom;
    cg.crop_x0 = 0; cg.crop_y0 = 0;

    const bool contiguous_frames = frame_stride == row_stride * (size_t)h;
    int pending[2] = {0, 0}, pending_f0[2] = {0, 0};
    int chunk = 0;
    for (int f0 = 0; f0 < n; f0 += kChunkFrames, chunk++) {
        const int b = chunk & 1;
        const int nf = n - f0 < kChunkFrames ? n - f0 : kChunkFrames;
        cudaStream_t st = ctx->streams[b];
        if (pending[b]) {       // results of the chunk that used this buffer two rounds ago
            MELF_CUDA(ctx, cudaStreamSynchronize(st));
            std::memcpy(out + pending_f0[b], ctx->h_res + (size_t)b * kChunkFrames,
                        sizeof(melf_result) * pending[b]);
            pending[b] = 0;
        }
        if (contiguous_frames) {
            cudaMemcpy3DParms p{};
            p.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t *>(bgr) + (size_t)f0 * frame_stride,
                                           row_stride, (size_t)w * 3, (size_t)h);
            p.srcPos = make_cudaPos((size_t)geom.crop_x0 * 3, (size_t)geom.crop_y0, 0);
            p.dstPtr = make_cudaPitchedPtr(ctx->d_chunk[b], pitch, crop_row_bytes, (size_t)geom.crop_h);
            p.dstPos = make_cudaPos(0, 0, 0);
            p.extent = make_cudaExtent(crop_row_bytes, (size_t)geom.crop_h, (size_t)nf);
            p.kind = cudaMemcpyHostToDevice;
            MELF_CUDA(ctx, cudaMemcpy3DAsync(&p, st));
        } else {
            for (int f = 0; f < nf; f++) {
                const uint8_t *src = bgr + (size_t)(f0 + f) * frame_stride +
                                     (size_t)geom.crop_y0 * row_stride + (size_t)geom.crop_x0 * 3;
                MELF_CUDA(ctx, cudaMemcpy2DAsync(ctx->d_chunk[b] + (size_t)f * frame_bytes, pitch, src,
                                                 row_stride, crop_row_bytes, (size_t)geom.crop_h,
                                                 cudaMemcpyHostToDevice, st));
            }
        }
        rc = run_path(ctx, ctx->d_chunk[b], nf, geom.crop_h, geom.crop_w, pitch, frame_bytes, cg,
                      ctx->d_chunk_keys[b], ctx->d_res[b], nullptr, st, 1 + b);
        if (rc) return rc;
        MELF_CUDA(ctx, cudaMemcpyAsync(ctx->h_res + (size_t)b * kChunkFrames, ctx->d_res[b],
                                       sizeof(melf_result) * nf, cudaMemcpyDeviceToHost, st));
        pending[b] = nf; pending_f0[b] = f0;
    }
    for (int b = 0; b < 2; b++)
        if (pending[b]) {
            MELF_CUDA(ctx, cudaStreamSynchronize(ctx->streams[b]));
            std::memcpy(out + pending_f0[b], ctx->h_res + (size_t)b * kChunkFrames,
                        sizeof(melf_result) * pending[b]);
        }
    return MELF_OK;
}

// ---- JPEG ingest ----
int melf_jpeg_info(const uint8_t *jpeg, size_t len, int *h, int *w, int *components, int *h_samp, int *v_samp)
{
    std::string err;
    if (!jpeg) return fail(nullptr, MELF_ERR_INVALID, "null stream");
    int rc = melf_jpeg_probe(jpeg, len, h, w, components, h_samp, v_samp, &err);
    return rc ? fail(nullptr, rc, err) : MELF_OK;
}

int melf_decode_jpeg(melf_ctx *ctx, const uint8_t *jpeg, const uint64_t *offsets, int n, int h, int w,
                     uint8_t *d_bgr, size_t row_stride, size_t frame_stride, void *stream)
{
    if (!ctx) return MELF_ERR_INVALID;
    if (n < 0 || h < 1 || w < 1 || (n > 0 && (!jpeg || !offsets || !d_bgr)))
        return fail(ctx, MELF_ERR_INVALID, "bad batch arguments");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    MelfJpegBatch batch{jpeg, offsets, n};
    std::string err;
    uint64_t launches = 0;
    JpegSpans spans(ctx);
    int rc = melf_jpeg_decode_range(&ctx->jpeg[0], batch, 0, n, h, w, 0, 0, w, h, d_bgr, row_stride,
                                    frame_stride, (cudaStream_t)stream, &err, &launches, spans.get());
    spans.commit(ctx);
    g_launches += launches;
    return rc ? fail(ctx, rc, err) : MELF_OK;
}

int melf_read_jpeg_batch(melf_ctx *ctx, const uint8_t *jpeg, const uint64_t *offsets, int n, melf_result *out)
{
    if (!ctx) return MELF_ERR_INVALID;
    if (n < 0 || (n > 0 && (!jpeg || !offsets || !out))) return fail(ctx, MELF_ERR_INVALID, "bad batch arguments");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    int h = 0, w = 0;
    std::string err;
    int rc = melf_jpeg_probe(jpeg + offsets[0], (size_t)(offsets[1] - offsets[0]), &h, &w, nullptr, nullptr,
                             nullptr, &err);
    if (rc) return fail(ctx, rc, "stream 0: " + err);
    MelfGeom geom;
    rc = make_geom(ctx, h, w, true, &geom);
    if (rc) return rc;
    // only the meter_rect pixels are converted to BGR: compact device layout [frame][crop_h][pitch]
    // (the rectangle starts at a column that is a multiple of 4, so that the colour kernel's
    // 12-byte pixel groups are aligned stores; the path is told where the real crop begins)
    const int jx0 = geom.crop_x0 & ~3, jw = geom.crop_w + (geom.crop_x0 - jx0);
    const size_t pitch = ((size_t)jw * 3 + 15) & ~(size_t)15;
    const size_t frame_bytes = pitch * geom.crop_h;
    // chunks in flight and streams per chunk (MELF_JPEG_SLOTS / MELF_JPEG_CHUNK override the defaults):
    // the batch is cut so that every slot is busy, in multiples of 512 streams
    static const int slots = [] {
        const char *e = std::getenv("MELF_JPEG_SLOTS");
        const int v = e ? std::atoi(e) : kJpegSlotsDefault;  // measured: profiles/r01_v11_jpeg_sweep.txt, r02_huffman_variants.txt
        return v < 1 ? 1 : v > kJpegSlots ? kJpegSlots : v;
    }();
    static const int chunk_cap = [] {
        const char *e = std::getenv("MELF_JPEG_CHUNK");
        const int v = e ? std::atoi(e) : kJpegChunkFrames;
        return v < 512 ? 512 : v > kJpegChunkFrames ? kJpegChunkFrames : v;
    }();
    static const bool trace = std::getenv("MELF_JPEG_TRACE") != nullptr;
    int chunk_frames = (n + slots - 1) / slots;
    chunk_frames = (chunk_frames + 511) & ~511;
    if (chunk_frames > chunk_cap) chunk_frames = chunk_cap;
    const size_t need = frame_bytes * chunk_frames;
    if (!ctx->h_jres) {
        for (int i = 0; i < slots; i++) {
            MELF_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->jstreams[i], cudaStreamNonBlocking));
            MELF_CUDA(ctx, cudaMalloc(&ctx->d_jres[i], sizeof(melf_result) * kJpegChunkFrames));
            MELF_CUDA(ctx, cudaMalloc(&ctx->d_jkeys[i], sizeof(u64) * kJpegChunkFrames));
        }
        MELF_CUDA(ctx, cudaMallocHost(&ctx->h_jres, sizeof(melf_result) * kJpegChunkFrames * kJpegSlots));
        MELF_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->jcopy, cudaStreamNonBlocking));
    }
    // all uploads of a batch on ONE stream, in chunk order (MELF_JPEG_COPY_STREAM=0: on the chunks' streams, A/B runs)
    static const bool own_copy_stream = !(std::getenv("MELF_JPEG_COPY_STREAM") && !std::atoi(std::getenv("MELF_JPEG_COPY_STREAM")));
    const cudaStream_t copy_stream = own_copy_stream ? ctx->jcopy : nullptr;
    if (need > ctx->jchunk_bytes) {
        for (int i = 0; i < slots; i++) {
            if (ctx->d_jchunk[i]) cudaFree(ctx->d_jchunk[i]);
            ctx->d_jchunk[i] = nullptr;
        }
        ctx->jchunk_bytes = 0;
        for (int i = 0; i < slots; i++) MELF_CUDA(ctx, cudaMalloc(&ctx->d_jchunk[i], need));
        ctx->jchunk_bytes = need;
    }
    MelfGeom cg = geom;
    cg.crop_x0 = geom.crop_x0 - jx0; cg.crop_y0 = 0;
    MelfJpegBatch batch{jpeg, offsets, n};
    int pending[kJpegSlots] = {}, pending_f0[kJpegSlots] = {};
    int chunk = 0;
    double host_ms = 0, wait_ms = 0;
    const auto t_call = std::chrono::steady_clock::now();
    auto ms_since = [](std::chrono::steady_clock::time_point t) {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t).count();
    };
    for (int f0 = 0; f0 < n; f0 += chunk_frames, chunk++) {
        const int b = chunk % slots;
        const int nf = n - f0 < chunk_frames ? n - f0 : chunk_frames;
        cudaStream_t st = ctx->jstreams[b];
        if (pending[b]) {
            const auto t0 = std::chrono::steady_clock::now();
            MELF_CUDA(ctx, cudaStreamSynchronize(st));
            std::memcpy(out + pending_f0[b], ctx->h_jres + (size_t)b * kJpegChunkFrames, sizeof(melf_result) * pending[b]);
            pending[b] = 0;
            wait_ms += ms_since(t0);
        }
        uint64_t launches = 0;
        JpegSpans spans(ctx);
        const auto t_host = std::chrono::steady_clock::now();
        rc = melf_jpeg_decode_range(&ctx->jpeg[1 + b], batch, f0, nf, h, w, jx0, geom.crop_y0,
                                    jw, geom.crop_h, ctx->d_jchunk[b], pitch, frame_bytes, st, &err,
                                    &launches, spans.get(), copy_stream);
        spans.commit(ctx);
        host_ms += ms_since(t_host);
        g_launches += launches;
        if (rc) return fail(ctx, rc, err);
        rc = run_path(ctx, ctx->d_jchunk[b], nf, geom.crop_h, jw, pitch, frame_bytes, cg,
                      ctx->d_jkeys[b], ctx->d_jres[b], nullptr, st, 3 + b);
        if (rc) return rc;
        MELF_CUDA(ctx, cudaMemcpyAsync(ctx->h_jres + (size_t)b * kJpegChunkFrames, ctx->d_jres[b],
                                       sizeof(melf_result) * nf, cudaMemcpyDeviceToHost, st));
        pending[b] = nf; pending_f0[b] = f0;
    }
    const double issued_ms = ms_since(t_call);
    for (int b = 0; b < slots; b++)
        if (pending[b]) {
            MELF_CUDA(ctx, cudaStreamSynchronize(ctx->jstreams[b]));
            std::memcpy(out + pending_f0[b], ctx->h_jres + (size_t)b * kJpegChunkFrames, sizeof(melf_result) * pending[b]);
        }
    if (trace)
        std::fprintf(stderr, "melf_read_jpeg_batch: n %d, %d chunks of %d in %d slots; header walk + launches %.2f ms, "
                     "slot waits %.2f ms, all issued at %.2f ms, done at %.2f ms\n",
                     n, chunk, chunk_frames, slots, host_ms, wait_ms, issued_ms, ms_since(t_call));
    return MELF_OK;
}

int melf_convert_hls(int device, const uint8_t *d_bgr, int n, int h, int w,
                     size_t row_stride, size_t frame_stride, int hue_shift, int s_tail,
                     uint8_t *d_hls, uint8_t *d_light,
                     const uint8_t *lo, const uint8_t *hi, uint8_t *d_mask, void *stream)
{
    if (n < 0 || h < 1 || w < 1 || (n > 0 && !d_bgr) || row_stride < (size_t)w * 3 ||
        hue_shift < 0 || hue_shift > 255 || (d_mask && (!lo || !hi)))
        return fail(nullptr, MELF_ERR_INVALID, "bad melf_convert_hls arguments");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(device);
    cudaError_t e = dev_scope.status;
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    MelfConvertArgs a{};
    a.bgr = d_bgr; a.row_stride = row_stride; a.frame_stride = frame_stride;
    a.n = n; a.h = h; a.w = w;
    a.hue_shift = hue_shift; a.s_tail = s_tail;
    a.hls = d_hls; a.light = d_light; a.mask = d_mask;
    a.lo = lo ? (lo[0] | (lo[1] << 8) | (lo[2] << 16)) : 0;
    a.hi = hi ? (hi[0] | (hi[1] << 8) | (hi[2] << 16)) : 0xffffff;
    e = melf_launch_convert(a, sms, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "melf_convert kernel launch");
    g_launches++;
    return MELF_OK;
}

int melf_in_range(int device, const uint8_t *d_img, int n, int h, int w, size_t row_stride, size_t frame_stride,
                  const uint8_t *lo, const uint8_t *hi, uint8_t *d_mask, void *stream)
{
    if (n < 0 || h < 1 || w < 1 || (n > 0 && (!d_img || !d_mask)) || !lo || !hi || row_stride < (size_t)w * 3)
        return fail(nullptr, MELF_ERR_INVALID, "bad melf_in_range arguments");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(device);
    if (dev_scope.status != cudaSuccess) return cuda_fail(nullptr, dev_scope.status, "cudaSetDevice");
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const cudaError_t e = melf_launch_in_range(d_img, row_stride, frame_stride, n, h, w,
                                               lo[0] | (lo[1] << 8) | (lo[2] << 16), hi[0] | (hi[1] << 8) | (hi[2] << 16),
                                               d_mask, sms, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "melf_in_range kernel launch");
    g_launches++;
    return MELF_OK;
}

int melf_match_template(melf_ctx *ctx, const uint8_t *d_img, int channels, int n, int h, int w,
                        size_t row_stride, size_t frame_stride, int use_rect,
                        int32_t *d_max_xy, float *d_max_val, float *d_result, void *stream)
{
    if (!ctx) return MELF_ERR_INVALID;
    if ((channels != 1 && channels != 3) || n < 0 || h < 1 || w < 1 || (n > 0 && !d_img) ||
        !d_max_xy || !d_max_val || row_stride < (size_t)w * channels)
        return fail(ctx, MELF_ERR_INVALID, "bad melf_match_template arguments");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    MelfGeom geom;
    int rc = make_geom(ctx, h, w, use_rect != 0, &geom);
    if (rc) return rc;
    rc = ensure_keys(ctx, n);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    rc = launch_locate(ctx, d_img, channels, n, h, w, row_stride, frame_stride, geom, ctx->d_keys,
                       d_result, st, 0);
    if (rc) return rc;
    MELF_CUDA(ctx, melf_launch_match_unpack(ctx->d_keys, n, geom.res_w, d_max_xy, d_max_val, st));
    g_launches++;
    return MELF_OK;
}

// melf_match_template without any allocation (include/meterelf_b200.h): the tensor-core locate kernel on the
// static scratch array.  Nothing here may call cudaMalloc / cudaFree / cudaEventCreate or a library that does.
int melf_match_template_noalloc(melf_ctx *ctx, const uint8_t *d_img, int channels, int n, int h, int w,
                                size_t row_stride, size_t frame_stride, int use_rect,
                                int32_t *d_max_xy, float *d_max_val, void *stream)
{
    if (!ctx) return MELF_ERR_INVALID;
    if ((channels != 1 && channels != 3) || n < 0 || h < 1 || w < 1 || (n > 0 && !d_img) ||
        !d_max_xy || !d_max_val || row_stride < (size_t)w * channels)
        return fail(ctx, MELF_ERR_INVALID, "bad melf_match_template_noalloc arguments");
    if (n > MELF_NOALLOC_MAX_FRAMES)
        return fail(ctx, MELF_ERR_UNSUPPORTED, "melf_match_template_noalloc: n = " + std::to_string(n) +
                    " frames exceeds MELF_NOALLOC_MAX_FRAMES = " + std::to_string(MELF_NOALLOC_MAX_FRAMES) +
                    " (static scratch); split the batch or use melf_match_template");
    if (!ctx->d_toeplitz || !melf_match_tc_supported(ctx->params.tmpl_w, ctx->params.tmpl_h))
        return fail(ctx, MELF_ERR_UNSUPPORTED, "melf_match_template_noalloc: the template does not fit the "
                    "tensor-core locate kernel (at most 193 x 128 px); use melf_match_template");
    if (n == 0) return MELF_OK;
    DeviceScope dev_scope(ctx->device);
    MELF_CUDA(ctx, dev_scope.status);
    MelfGeom geom;
    int rc = make_geom(ctx, h, w, use_rect != 0, &geom);
    if (rc) return rc;
    unsigned char *scratch = nullptr;
    MELF_CUDA(ctx, noalloc_setup(ctx->device, &scratch));          // normally done by melf_create already
    const int ctas = ctx->num_sms < kNoAllocCtas ? ctx->num_sms : kNoAllocCtas;
    if (!scratch || melf_match_fused_ov_scratch_bytes(ctas) > kNoAllocKeysOff)
        return fail(ctx, MELF_ERR_UNSUPPORTED, "melf_match_template_noalloc: static scratch unavailable or too small");
    cudaStream_t st = (cudaStream_t)stream;
    u64 *keys = reinterpret_cast<u64 *>(scratch + kNoAllocKeysOff);
    // every call shares the array and it keeps the last call's contents: the keys must start at zero (atomicMax
    // slots), and the slab rows a geometry never writes must be zero whatever an earlier geometry left there
    MELF_CUDA(ctx, cudaMemsetAsync(scratch, 0, melf_match_fused_ov_scratch_bytes(ctas), st));
    MELF_CUDA(ctx, cudaMemsetAsync(keys, 0, (size_t)n * sizeof(u64), st));
    MelfMatchTcArgs t{};
    t.img = d_img;
    const uint8_t *end = d_img + (size_t)(n - 1) * frame_stride + (size_t)(h - 1) * row_stride + (size_t)w * channels;
    t.img_limit = reinterpret_cast<const uint8_t *>((reinterpret_cast<uintptr_t>(end) + 15) & ~(uintptr_t)15);
    t.row_stride = row_stride; t.frame_stride = frame_stride;
    t.channels = channels;
    t.img_w = w; t.img_h = h; t.img_n = n;
    t.geom = geom;
    t.toeplitz = ctx->d_toeplitz;
    t.tmpl_w = ctx->params.tmpl_w; t.tmpl_h = ctx->params.tmpl_h;
    t.sum_t = ctx->sum_t; t.n_t = ctx->n_t;
    t.match_key = keys;
    t.result = nullptr;
    melf_match_tc_plan(t, n);
    t.slabs = scratch;
    t.max_items = t.n_items;
    MELF_CUDA(ctx, melf_launch_match_fused_ov(t, ctas, st));
    MELF_CUDA(ctx, melf_launch_match_unpack(keys, n, geom.res_w, d_max_xy, d_max_val, st));
    g_launches += 2;
    return MELF_OK;
}

}  // extern "C"
