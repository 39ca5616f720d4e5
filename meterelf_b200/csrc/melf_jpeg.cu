// Ingest row: a batch of baseline JPEG streams -> BGR frames on the device,
// bit-identical to what cv2.imread returns (reference call site meterelf/_image.py:49; the decoder
// behind it is OpenCV's bundled libjpeg-turbo with libjpeg defaults: ISLOW IDCT, fancy
// upsampling, BGR output -- restated on the CPU in oracle/jpeg_oracle.c).
//
// The host only walks the marker segments (a few hundred bytes per stream) and uploads the
// compressed bytes; everything else runs on the GPU:
//
//   melf_jpeg_destuff_kernel   one CTA per stream: removes the FF00 byte stuffing of the entropy
//                              segment and cuts it at the first marker (two passes: first marker,
//                              then a block-wide scan compaction).
//   melf_jpeg_huffman_kernel   sequential Huffman decoding, one lane per stream (G lanes of a
//                              one-warp CTA active, so that a batch spreads over all SMs); 10-bit
//                              look-ahead tables built in shared memory straight from the DHT bytes
//                              of the uploaded stream; coefficients stored in zig-zag order.
//   melf_jpeg_idct_kernel      one thread per 8x8 block: dequantisation with the DQT bytes of the
//                              stream + the 13-bit integer LLM inverse DCT (jidctint.c arithmetic),
//                              all in registers, samples written to per-component planes.
//   melf_jpeg_color_kernel     triangle-filter ("fancy") chroma upsampling + YCbCr -> BGR for the
//                              requested rectangle of every frame, 4 pixels per thread.
//
// All four are byte/integer kernels bounded by latency (Huffman) or HBM (the rest); no tensor work.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "melf_jpeg.h"

namespace {

#ifndef MELF_JPEG_LOOK
#define MELF_JPEG_LOOK 10
#endif
constexpr int kLook = MELF_JPEG_LOOK;    // look-ahead bits of the Huffman tables in shared memory
constexpr int kLutEntries = 1 << kLook;

struct JpegDesc {                        // one per stream (device copy)
    uint64_t base;                       // offset of the stream in the blob
    uint64_t clean_off;                  // offset of its de-stuffed entropy bytes in the clean buffer
    uint32_t scan_off, scan_len;         // entropy-coded segment, relative to base
    uint32_t huff_off[4];                // luma DC, luma AC, chroma DC, chroma AC: the 16 counts, then the symbols
    uint32_t quant_off[3];               // 64 zig-zag entries per component ...
    uint32_t quant16;                    // ... bit c set: 16-bit big-endian entries
    uint32_t restart;                    // restart interval in MCUs (0: none)
};

struct JpegGeom {                        // shared by every stream of a batch
    int w, h, ncomp, hs, vs;             // luma sampling factors (chroma is 1x1)
    int mcus_x, mcus_y, nslots;          // blocks per MCU
    uint32_t slot_info;                  // 4 bits per slot: comp | bx << 2 | by << 3
    int blocks_w[3], blocks_h[3];        // MCU-padded block grid per component
    int block_off[3];                    // first block of the component inside a stream's coefficient run
    int blocks_per_image;
    int plane_off[3];                    // byte offset of the component's sample plane inside a stream's planes
    int plane_bytes;
    int ds_w, ds_h;                      // downsampled_width / height of the chroma components
};

// The MCUs whose blocks the later stages read (the output rectangle plus the chroma apron).  The
// Huffman stage keeps only their coefficients, densely, in decode order, stops after the last of them,
// and sends every other block to one trash block at the end of the stream's run.
struct McuRect {
    int mx0, mx1, my0, my1;              // inclusive MCU range
    int mw;                              // mx1 - mx0 + 1
    int stride_blocks;                   // kept blocks per stream + 1 (the trash block)
};

// ---------------------------------------------------------------------------------------------
// host: marker walk (ITU-T T.81 Annex B).  Fills `d` and checks the geometry against `g`
// (the first stream defines it).  Returns 0, or an error text.
// ---------------------------------------------------------------------------------------------
int rd16(const uint8_t *p) { return (p[0] << 8) | p[1]; }

const char *check_huffman(const uint8_t *seg, int avail, int *used)
{
    if (avail < 17) return "truncated DHT";
    int total = 0;
    unsigned code = 0;
    for (int len = 1; len <= 16; len++) {
        total += seg[len];
        code += seg[len];
        if (code > (1u << len)) return "over-subscribed Huffman table";
        code <<= 1;
    }
    if (total > 256 || 17 + total > avail) return "truncated DHT";
    *used = 17 + total;
    return nullptr;
}

const char *parse_stream(const uint8_t *data, size_t len, JpegDesc *d, JpegGeom *g, bool first)
{
    if (len < 4 || data[0] != 0xFF || data[1] != 0xD8) return "not a JPEG stream";
    uint32_t dht[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}}, dqt[4] = {0, 0, 0, 0};
    bool dqt16[4] = {false, false, false, false};
    int w = 0, h = 0, ncomp = 0, chs[3] = {0, 0, 0}, cvs[3] = {0, 0, 0}, ctq[3] = {0, 0, 0}, cid[3] = {0, 0, 0};
    bool sof = false;
    bool saw_jfif = false, saw_adobe = false;       // libjpeg jdmarker.c examine_app0 / examine_app14
    int adobe_transform = 0;
    size_t i = 2;
    for (;;) {
        if (i + 4 > len || data[i] != 0xFF) return "damaged JPEG headers";
        while (i < len && data[i] == 0xFF) i++;
        if (i >= len) return "damaged JPEG headers";
        const int m = data[i++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (m == 0xD9 || i + 2 > len) return "damaged JPEG headers";
        const int L = rd16(data + i);
        if (L < 2 || i + (size_t)L > len) return "damaged JPEG headers";
        const uint8_t *seg = data + i + 2;
        const int n = L - 2;
        if (m == 0xDB) {
            int o = 0;
            while (o < n) {
                const int pq = seg[o] >> 4, tq = seg[o] & 15;
                o++;
                if (tq > 3 || pq > 1 || o + 64 * (pq + 1) > n) return "damaged DQT";
                dqt[tq] = (uint32_t)(seg + o - data);
                dqt16[tq] = pq != 0;
                o += 64 * (pq + 1);
            }
        } else if (m == 0xC4) {
            int o = 0;
            while (o < n) {
                const int tc = seg[o] >> 4, th = seg[o] & 15;
                if (tc > 1 || th > 3) return "damaged DHT";
                int used = 0;
                const char *e = check_huffman(seg + o, n - o, &used);
                if (e) return e;
                dht[tc][th] = (uint32_t)(seg + o + 1 - data);
                o += used;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (n < 6 || sof) return "damaged SOF";
            if (seg[0] != 8) return "unsupported: sample precision is not 8 bits";
            h = rd16(seg + 1);
            w = rd16(seg + 3);
            ncomp = seg[5];
            if (ncomp != 1 && ncomp != 3) return "unsupported: not a 1- or 3-component JPEG";
            if (n < 6 + 3 * ncomp) return "damaged SOF";
            for (int c = 0; c < ncomp; c++) {
                cid[c] = seg[6 + 3 * c];
                chs[c] = seg[7 + 3 * c] >> 4;
                cvs[c] = seg[7 + 3 * c] & 15;
                ctq[c] = seg[8 + 3 * c];
                if (ctq[c] > 3) return "damaged SOF";
            }
            sof = true;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC8 && m != 0xCC) {
            return "unsupported: progressive, lossless or arithmetic-coded JPEG";
        } else if (m == 0xDD) {
            if (n < 2) return "damaged DRI";
            d->restart = (uint32_t)rd16(seg);
        } else if (m == 0xE0) {
            if (n >= 14 && !std::memcmp(seg, "JFIF\0", 5)) saw_jfif = true;
        } else if (m == 0xEE) {
            if (n >= 12 && !std::memcmp(seg, "Adobe", 5)) { saw_adobe = true; adobe_transform = seg[11]; }
        } else if (m == 0xDA) {
            if (!sof) return "SOS before SOF";
            if (n < 1 || seg[0] != ncomp || n < 1 + 2 * ncomp + 3) return "unsupported: multi-scan JPEG";
            int td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
            for (int c = 0; c < ncomp; c++) {
                if (seg[1 + 2 * c] != cid[c]) return "unsupported: scan component order";
                td[c] = seg[2 + 2 * c] >> 4;
                ta[c] = seg[2 + 2 * c] & 15;
                if (td[c] > 3 || ta[c] > 3) return "damaged SOS";
            }
            if (ncomp == 3 && (td[1] != td[2] || ta[1] != ta[2]))
                return "unsupported: the chroma components use different Huffman tables";
            const int cc = ncomp == 3 ? 1 : 0;
            d->huff_off[0] = dht[0][td[0]];
            d->huff_off[1] = dht[1][ta[0]];
            d->huff_off[2] = dht[0][td[cc]];
            d->huff_off[3] = dht[1][ta[cc]];
            for (int k = 0; k < 4; k++)
                if (!d->huff_off[k]) return "missing Huffman table";
            d->quant16 = 0;
            for (int c = 0; c < 3; c++) {
                const int cu = c < ncomp ? c : 0;
                if (!dqt[ctq[cu]]) return "missing quantisation table";
                d->quant_off[c] = dqt[ctq[cu]];
                if (dqt16[ctq[cu]]) d->quant16 |= 1u << c;
            }
            d->scan_off = (uint32_t)(i + (size_t)L);
            if (len - (i + (size_t)L) > 0x0FFFFFFFu) return "unsupported: entropy-coded segment larger than 256 MB";
            d->scan_len = (uint32_t)(len - (i + (size_t)L));
            break;
        }
        i += (size_t)L;
    }
    int hs = 1, vs = 1;
    if (ncomp == 3) {
        // libjpeg jdapimin.c default_decompress_parms: JFIF means YCbCr; else an Adobe marker decides
        // (transform 0 = RGB); else the ids 'R','G','B' mean RGB.  The kernels convert YCbCr only, so
        // RGB files go back to the caller (the Python layer hands them to cv2.imread).
        const bool rgb = saw_jfif ? false : saw_adobe ? adobe_transform == 0
                                                      : (cid[0] == 82 && cid[1] == 71 && cid[2] == 66);
        if (rgb) return "unsupported: RGB colour space (Adobe transform 0 or component ids R, G, B)";
        hs = chs[0];
        vs = cvs[0];
        if (!((hs == 1 && vs == 1) || (hs == 2 && vs == 1) || (hs == 2 && vs == 2)) ||
            chs[1] != 1 || cvs[1] != 1 || chs[2] != 1 || cvs[2] != 1)
            return "unsupported: chroma subsampling other than 4:4:4, 4:2:2 or 4:2:0";
    }
    if (first) {
        std::memset(g, 0, sizeof *g);
        g->w = w; g->h = h; g->ncomp = ncomp; g->hs = hs; g->vs = vs;
        g->mcus_x = (w + 8 * hs - 1) / (8 * hs);
        g->mcus_y = (h + 8 * vs - 1) / (8 * vs);
        int slot = 0;
        for (int c = 0; c < ncomp; c++) {
            const int ch = c ? 1 : hs, cv = c ? 1 : vs;
            for (int by = 0; by < cv; by++)
                for (int bx = 0; bx < ch; bx++, slot++) g->slot_info |= (uint32_t)(c | bx << 2 | by << 3) << (4 * slot);
            g->blocks_w[c] = g->mcus_x * ch;
            g->blocks_h[c] = g->mcus_y * cv;
            g->block_off[c] = g->blocks_per_image;
            g->blocks_per_image += g->blocks_w[c] * g->blocks_h[c];
            g->plane_off[c] = g->plane_bytes;
            g->plane_bytes += g->blocks_w[c] * g->blocks_h[c] * 64;
        }
        g->nslots = slot;
        g->ds_w = (w + hs - 1) / hs;
        g->ds_h = (h + vs - 1) / vs;
    } else if (w != g->w || h != g->h || ncomp != g->ncomp || hs != g->hs || vs != g->vs) {
        return "the streams of a batch must share size, component count and subsampling";
    }
    return nullptr;
}

// ---------------------------------------------------------------------------------------------
// kernel 0: remove byte stuffing, cut at the first marker
// ---------------------------------------------------------------------------------------------
constexpr int DESTUFF_THREADS = 128;
constexpr int DESTUFF_PER_THREAD = 16;

__global__ void __launch_bounds__(DESTUFF_THREADS)
melf_jpeg_destuff_kernel(const uint8_t *__restrict__ blob, const JpegDesc *__restrict__ desc,
                         uint8_t *__restrict__ clean, uint32_t *__restrict__ clean_len)
{
    __shared__ uint32_t s_first;
    __shared__ uint32_t s_warp[DESTUFF_THREADS / 32];
    __shared__ uint32_t s_base;
    const JpegDesc d = desc[blockIdx.x];
    const uint8_t *src = blob + d.base + d.scan_off;
    const uint32_t len = d.scan_len;
    const int tid = threadIdx.x;
    if (tid == 0) { s_first = len; s_base = 0; }
    __syncthreads();
    // pass 1: first 0xFF that is followed by neither 0x00 nor a restart marker code (the end of the data)
    // (16 aligned bytes per load; the bytes in front of the segment and behind it belong to the blob
    // allocation -- headers in front, 16 bytes of slack behind -- and are masked out by the index test;
    // only the ~6 % of vectors that hold a 0xFF byte are looked at byte by byte)
    uint32_t first = len;
    {
        const uint32_t a0 = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15);
        const uint4 *vsrc = reinterpret_cast<const uint4 *>(src - a0);
        const uint32_t nvec = (a0 + len + 15) >> 4;
        for (uint32_t v = tid; v < nvec && first == len; v += DESTUFF_THREADS) {
            const uint4 q = __ldg(vsrc + v);
            const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
            uint32_t any = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t inv = ~w4[k];                          // a 0xFF byte becomes a zero byte
                any |= (inv - 0x01010101u) & ~inv & 0x80808080u;
            }
            if (any == 0) continue;
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const int64_t j = (int64_t)v * 16 + k - a0;
                if (j < 0 || j >= (int64_t)len || (w4[k >> 2] >> (8 * (k & 3)) & 255) != 0xFF) continue;
                const uint8_t nx = j + 1 < (int64_t)len ? src[j + 1] : 0xD9;
                if (nx != 0x00 && (nx & 0xF8) != 0xD0) { first = (uint32_t)j; break; }
            }
        }
    }
    if (first < len) atomicMin(&s_first, first);
    __syncthreads();
    const uint32_t end = s_first;
    // pass 2: copy [0, end) without the zero that follows a data 0xFF and without the RSTn markers
    // (whole bytes, so the decoder's byte alignment at a restart survives)
    uint8_t *dst = clean + d.clean_off;
    const uint32_t tile = DESTUFF_THREADS * DESTUFF_PER_THREAD;
    for (uint32_t t0 = 0; t0 < end; t0 += tile) {
        const uint32_t j0 = t0 + tid * DESTUFF_PER_THREAD;
        uint8_t b[DESTUFF_PER_THREAD];
        uint32_t keep = 0;
        uint8_t prev = (j0 > 0 && j0 <= end) ? src[j0 - 1] : 0;
#pragma unroll
        for (int k = 0; k < DESTUFF_PER_THREAD; k++) {
            const uint32_t j = j0 + k;
            b[k] = j < end ? src[j] : 0;
            const uint8_t nx = (b[k] == 0xFF && j + 1 < end) ? src[j + 1] : 0;
            const bool drop = (prev == 0xFF && (b[k] == 0x00 || (b[k] & 0xF8) == 0xD0)) ||
                              (b[k] == 0xFF && (nx & 0xF8) == 0xD0);
            if (j < end && !drop) keep |= 1u << k;
            prev = (prev == 0xFF && b[k] == 0x00) ? 0x01 : b[k];   // a stuffed zero ends the FF pair
        }
        const uint32_t cnt = __popc(keep);
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if ((tid & 31) >= o) incl += v;
        }
        if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
        __syncthreads();
        uint32_t before = s_base, total = 0;
#pragma unroll
        for (int wi = 0; wi < DESTUFF_THREADS / 32; wi++) {
            if (wi < (tid >> 5)) before += s_warp[wi];
            total += s_warp[wi];
        }
        uint32_t o = before + incl - cnt;
#pragma unroll
        for (int k = 0; k < DESTUFF_PER_THREAD; k++)
            if (keep >> k & 1) dst[o++] = b[k];
        __syncthreads();
        if (tid == 0) s_base += total;
        __syncthreads();
    }
    // zero padding: the decoder reads whole words and is fed zero bits after the data (jdhuff.c)
    const uint32_t out_len = s_base;
    for (uint32_t j = out_len + tid; j < ((out_len + 15) & ~15u) + 16; j += DESTUFF_THREADS) dst[j] = 0;
    if (tid == 0) clean_len[blockIdx.x] = out_len;
}

// ---------------------------------------------------------------------------------------------
// kernel 1: Huffman decoding (T.81 F.2.2; libjpeg jdhuff.c decode_mcu), one lane per stream
// ---------------------------------------------------------------------------------------------
struct HuffSmem {                         // per stream
    uint16_t lut[4][kLutEntries];         // (code length << 8) | symbol for codes of <= kLook bits, else 0
    uint32_t limit[4][8];                 // lengths kLook+1 .. 16: first 16-bit window value past the codes of
                                          // that length (canonical codes are contiguous, so monotone)
    int32_t valoff[4][8];                 // ... and symbol index - code of the first code of that length
    uint8_t vals[4][256];                 // the table's symbol bytes (codes longer than kLook bits look them up
                                          // here: with 16 streams per warp some lane does so in most iterations,
                                          // and a global load there was the kernel's top stall, ncu long_scoreboard)
};

// Builds the tables of one stream with the whole warp.
__device__ void build_tables(HuffSmem &S, const uint8_t *__restrict__ stream, const JpegDesc &d, int lane)
{
#pragma unroll 1
    for (int t = 0; t < 4; t++) {
        const uint32_t hoff = t == 0 ? d.huff_off[0] : t == 1 ? d.huff_off[1] : t == 2 ? d.huff_off[2] : d.huff_off[3];
        const uint8_t *counts = stream + hoff;
        const uint8_t *symbols = counts + 16;
        // canonical code assignment (T.81 Annex C), the same on every lane
        int first[17], valptr[17], cnt[17];
        int code = 0, k = 0;
#pragma unroll
        for (int l = 1; l <= 16; l++) {
            cnt[l] = counts[l - 1];
            first[l] = code;
            valptr[l] = k;
            code = (code + cnt[l]) << 1;
            k += cnt[l];
        }
        for (int i = lane; i < kLutEntries; i += 32) {
            uint16_t e = 0;
#pragma unroll
            for (int l = 1; l <= kLook; l++) {
                const int c = i >> (kLook - l);
                if (e == 0 && c >= first[l] && c - first[l] < cnt[l])
                    e = (uint16_t)((l << 8) | symbols[valptr[l] + c - first[l]]);
            }
            S.lut[t][i] = e;
        }
        if (lane < 8) {
            uint32_t lim = 0xFFFFFFFFu;
            int vo = 0;
#pragma unroll
            for (int l = kLook + 1; l <= 16; l++)
                if (l - kLook - 1 == lane) {
                    lim = (uint32_t)(first[l] + cnt[l]) << (16 - l);
                    vo = valptr[l] - first[l];
                }
            S.limit[t][lane] = lim;
            S.valoff[t][lane] = vo;
        }
        for (int i = lane; i < 256; i += 32) S.vals[t][i] = i < k ? symbols[i] : 0;
    }
}

// The loop body is written with selects instead of branches: the lanes of a warp decode different
// streams and would otherwise serialise on every DC/AC/end-of-block/refill decision.  Coefficients
// go out in DECODE order (block b of the scan at out + 64*b, zig-zag order inside the block), so
// that the end of a block costs one pointer increment; the IDCT kernel undoes both orders.
// SHARED: every stream of the launch carries the same four Huffman tables (the host compared the
// DHT bytes), so one set of tables per CTA serves all lanes.
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr)
{
    uint16_t v;
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}

// MODE (round 2, the instruction count per symbol is what bounds this stage):
//   HUFF_CAREFUL  everything: restart intervals and jdhuff.c's behaviour when a stream runs out of data
//   HUFF_FAST     for chunks without restart intervals: none of that in the loop (the end-of-MCU path shrinks from
//                 ~60 to ~10 instructions, and some lane of the warp takes it in a third of the iterations).  A
//                 stream that read past its end is only FLAGGED (flags[stream] = 1, else 0) ...
//   HUFF_REDO     ... and decoded again by this careful variant, launched right behind on the same stream: warps
//                 without a flagged stream return at once, the others clear the coefficients of the flagged
//                 streams (the fast pass wrote rubbish behind the end of the data) and decode them as HUFF_CAREFUL.
enum { HUFF_CAREFUL = 0, HUFF_FAST = 1, HUFF_REDO = 2 };
#ifndef MELF_HUFF_LEAN
#define MELF_HUFF_LEAN 1     // 0: HUFF_FAST as the generic loop with its restart / out-of-data parts compiled out (A/B builds)
#endif

// The decoder consumes its stream one 32-bit word per ~5 symbols and holds one word in flight (nextw): a word that
// misses L1 comes from L2 / DRAM (the destuffed chunk does not fit L2) in 600+ clk, several times the ~300 clk
// between two refills -- 28 % of all stall samples of the lean loop sat on the first use of that word
// (profiles/r02_ncu_huffman_lean_source.txt).  So every eighth refill asks for the 32-byte sector one cache line
// further on; the request needs no register and nothing waits for it.
#ifndef MELF_HUFF_DEBUG
#define MELF_HUFF_DEBUG 0
#endif
#ifndef MELF_HUFF_PREFETCH
#define MELF_HUFF_PREFETCH 0     // measured: no gain, the lone launch 10.8 -> 11.7 ms (profiles/r02_huffman_variants.txt)
#endif
__device__ __forceinline__ void huff_prefetch(const uint32_t *src, int wp, int wmax, bool need)
{
#if MELF_HUFF_PREFETCH
    if (need && (wp & 7) == 0 && wp + 32 < wmax)
        asm volatile("prefetch.global.L1 [%0];" :: "l"(src + wp + 32));
#endif
}

// W warps per CTA (W > 1 only with SHARED tables and the lean loop): one table set serves W warps (8 by default), so
// that shared memory (9.5 KB per table set) no longer caps the resident warps at 24 per SM.
template <int G, bool SHARED, int MODE, int W = 1>
__global__ void __launch_bounds__(32 * W)
melf_jpeg_huffman_kernel(const uint8_t *__restrict__ blob, const uint8_t *__restrict__ clean,
                         const JpegDesc *__restrict__ desc, const uint32_t *__restrict__ order,
                         const uint32_t *__restrict__ clean_len, uint32_t *__restrict__ flags,
                         int n, const JpegGeom g, const McuRect R, int16_t *__restrict__ coef)
{
    extern __shared__ __align__(16) unsigned char huff_smem[];
    HuffSmem *S = reinterpret_cast<HuffSmem *>(huff_smem);
    const int lane = threadIdx.x & 31;
    const int img0 = (blockIdx.x * W + (threadIdx.x >> 5)) * G;
    if (W > 1) {
        // (SHARED tables, lean loop) warp 0 builds the CTA's table set from the CTA's first stream
        if (threadIdx.x < 32) {
            const JpegDesc d = desc[order[blockIdx.x * W * G]];
            build_tables(S[0], blob + d.base, d, lane);
        }
        __syncthreads();
        if (lane >= G || img0 + lane >= n) return;
    } else
    if (MODE == HUFF_REDO) {
        const bool mine = lane < G && img0 + lane < n && flags[order[img0 + lane]] != 0;
        uint32_t todo = __ballot_sync(0xffffffffu, mine);
        if (!todo) return;
        // the whole warp clears the coefficients of every flagged stream of this warp
        const size_t per_stream = (size_t)R.stride_blocks * 64;               // int16 elements, a multiple of 8 bytes
        while (todo) {
            const int l = __ffs(todo) - 1;
            todo &= todo - 1;
            uint2 *dst = reinterpret_cast<uint2 *>(coef + (size_t)order[img0 + l] * per_stream);
            for (size_t i = lane; i < per_stream / 4; i += 32) dst[i] = make_uint2(0u, 0u);
        }
        __syncwarp();
        for (int gi = 0; gi < (SHARED ? 1 : G) && img0 + gi < n; gi++) {
            const JpegDesc d = desc[order[img0 + gi]];
            build_tables(S[gi], blob + d.base, d, lane);
        }
        __syncwarp();
        if (!mine) return;
    } else {
        for (int gi = 0; gi < (SHARED ? 1 : G) && img0 + gi < n; gi++) {
            const JpegDesc d = desc[order[img0 + gi]];
            build_tables(S[gi], blob + d.base, d, lane);
        }
        __syncwarp();
        if (lane >= G || img0 + lane >= n) return;
    }
    const int img = (int)order[img0 + lane];            // lanes take the streams longest first (host-sorted)
    const HuffSmem &T = S[SHARED ? 0 : lane];
    // the look-ahead tables are read through a shared-window address computed once (a generic pointer
    // makes the compiler rebuild the window base in every iteration, on the critical path of the lookup)
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(&T.lut[0][0]);
    const int restart = (int)desc[img].restart;
#if MELF_HUFF_DEBUG & 1       // measurement only: every lane reads the stream of the warp's first lane
    const uint32_t *src = reinterpret_cast<const uint32_t *>(clean + desc[order[img0]].clean_off);
#else
    const uint32_t *src = reinterpret_cast<const uint32_t *>(clean + desc[img].clean_off);
#endif
    const int wmax = (int)((clean_len[img] + 3) >> 2);
    const uint32_t avail_bits = clean_len[img] * 8u;           // parse_stream caps a scan at 256 MB
    int16_t *kept = coef + (size_t)img * R.stride_blocks * 64;
    int16_t *const trash = kept + (size_t)(R.stride_blocks - 1) * 64;
    // kept / skipped MCUs alternate in runs along the scan: my0 * mcus_x + mx0 skipped ones first, then
    // mw kept and mcus_x - mw skipped ones per MCU row; `run` counts the current run down
    const int mcus_x = g.mcus_x;
    const int run_keep = R.mw == mcus_x ? 0x7fffffff : R.mw, run_skip = mcus_x - R.mw;
    int run = R.my0 * mcus_x + R.mx0;
    bool keep = run == 0;
    run = keep ? run_keep : run;
    int16_t *blk = keep ? kept : trash;

    // bit reader: w0 = the 32 bits being consumed (first bit of the stream in the MSB), w1 = the 32 bits
    // after them, bp = bits of w0 already used (< 32 once refilled), nextw = the word after w1 (as loaded).
    // A code (<= 16 bits) and its extra bits (<= 15) always fit the 32-bit window funnel-shifted out of w0:w1.
    uint32_t w0 = __byte_perm(wmax > 0 ? src[0] : 0, 0, 0x0123), w1 = __byte_perm(wmax > 1 ? src[1] : 0, 0, 0x0123);
    int wp = 2, bp = 0;
    uint32_t nextw = wp < wmax ? src[wp] : 0;
    // DC predictors: (pred, pred1, pred2) belong to components (comp, comp + 1, comp + 2) mod 3 and rotate
    // when the component changes at the end of a block, so the symbol loop touches only `pred`
    int pred = 0, pred1 = 0, pred2 = 0;
    int slot = 0, k = 0, comp = 0, in_interval = 0;
    const int nslots = g.nslots;
    const uint32_t slot_info = g.slot_info;
    int left = (R.my1 * mcus_x + R.mx1 + 1) * nslots;          // nothing after the last kept MCU is needed
    if (MODE == HUFF_FAST && MELF_HUFF_LEAN) {
        // The lean loop.  Same decoding as the loop below, minus restart / out-of-data handling, written for the
        // latency of ONE warp (466 clk and 82 instructions per symbol before, profiles/r02_ncu_huffman_fast_source.txt):
        // table bases of the current component kept in registers (no table index arithmetic per symbol), the shared
        // window base computed once by volatile asm (the compiler re-derived it from SR_CgaCtaId in every iteration),
        // 32-bit coefficient offsets instead of 64-bit pointer selects, sign extension without compares.
        uint32_t lut0;
        asm volatile("{\n\t.reg .u64 a;\n\tcvta.to.shared.u64 a, %1;\n\tcvt.u32.u64 %0, a;\n\t}"
                     : "=r"(lut0) : "l"(&T.lut[0][0]));
        constexpr uint32_t kTab = kLutEntries * 2;
        uint32_t lut_dc = lut0, lut_ac = lut0 + kTab;                  // component 0
        int16_t *const base = kept;
        const uint32_t trash_off = (uint32_t)(R.stride_blocks - 1) * 64u;
        uint32_t kept_off = 0, blk_off = keep ? 0u : trash_off;
        uint32_t nextw2 = src[min(wp + 1, wmax + 2)];
#if MELF_HUFF_DEBUG & 4       // measurement only: per-stream clocks and iterations instead of the redo flag
        const long long dbg_t0 = clock64();
        uint32_t dbg_iters = 0;
#endif
        while (left > 0) {
#if MELF_HUFF_DEBUG & 4
            ++dbg_iters;
#endif
            {
                const bool need = bp >= 32;
                bp -= need ? 32 : 0;
                w0 = need ? w1 : w0;
                w1 = need ? __byte_perm(nextw, 0, 0x0123) : w1;
                wp += need ? 1 : 0;
                // two words in flight (nextw = word wp, nextw2 = word wp + 1): the load issued here is used two
                // refills (~3000 clk) later; with one word in flight 28 % of all stall samples sat on its first use
                nextw = need ? nextw2 : nextw;
                const uint32_t *ptr = src + min(wp + 1, wmax + 2);     // >= 16 zero bytes follow the destuffed data
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\t@p ld.global.nc.u32 %0, [%1];\n\t}"
                             : "+r"(nextw2) : "l"(ptr), "r"((int)need));
            }
            const bool is_dc = k == 0;
            const uint32_t win = __funnelshift_l(w1, w0, bp);
            const uint32_t e = lds_u16((is_dc ? lut_dc : lut_ac) + ((win >> (32 - kLook)) << 1));
            int len = (int)(e >> 8), sym = (int)(e & 255);
            if (e == 0) {                                    // code longer than kLook bits (rare)
                const int t = (comp ? 2 : 0) + (is_dc ? 0 : 1);
                const uint32_t win16 = win >> 16;
                int l = kLook + 1;
#pragma unroll
                for (int j = 0; j < 16 - kLook - 1; j++) l += win16 >= T.limit[t][j] ? 1 : 0;
                const bool valid = win16 < T.limit[t][16 - kLook - 1];
                const int idx = (int)(win16 >> (16 - l)) + T.valoff[t][l - kLook - 1];
                sym = valid ? T.vals[t][idx & 255] : 0;
                len = l;
            }
            const int s = sym & 15, r = sym >> 4;
            const uint32_t field = win << len;                          // the s extra bits from bit 31 down
            const uint32_t bits = (field >> 1) >> (31 - s);
            bp += len + s;
            // EXTEND: a field whose first bit is 0 stands for bits - (2^s - 1); s = 0 gives 0 either way
            const int v = (int)bits + (~((int)field >> 31) & (1 - (1 << s)));
            pred += is_dc ? v : 0;
            const int pos = is_dc ? 0 : k + r;
            const bool st = is_dc || (s != 0 && pos < 64);
#if MELF_HUFF_DEBUG & 2       // measurement only: every store goes to the spare block
            base[trash_off + (uint32_t)(pos & 63) + (st ? 0u : 0u)] = (int16_t)(is_dc ? pred : v);
#else
            base[st ? blk_off + (uint32_t)pos : trash_off] = (int16_t)(is_dc ? pred : v);
#endif
            k = (is_dc || s != 0) ? pos + 1 : (r == 15 ? k + 16 : 64);
            if (k >= 64) {
                k = 0;
                kept_off += keep ? 64u : 0u;
                left--;
                const bool mcu_end = slot + 1 == nslots;
                slot = mcu_end ? 0 : slot + 1;
                const int ncomp = (int)(slot_info >> (4 * slot) & 3);
                if (ncomp != comp) {
                    const int t0 = pred;
                    pred = pred1; pred1 = pred2; pred2 = t0;
                    comp = ncomp;
                    lut_dc = lut0 + (comp ? 2 * kTab : 0u);
                    lut_ac = lut_dc + kTab;
                }
                if (mcu_end && --run == 0) {
                    keep = !keep;
                    run = keep ? run_keep : run_skip;
                }
                blk_off = keep ? kept_off : trash_off;
            }
        }
#if MELF_HUFF_DEBUG & 4
        flags[img] = ((uint32_t)((clock64() - dbg_t0) >> 9) << 16) | ((dbg_iters >> 2) & 0xffffu);
        return;
#endif
        flags[img] = (uint32_t)(wp - 2) * 32u + (uint32_t)bp > avail_bits ? 1u : 0u;
        return;
    }
    while (left > 0) {
        {   // refill: w0 is used up
            const bool need = bp >= 32;
            bp -= need ? 32 : 0;
            w0 = need ? w1 : w0;
            w1 = need ? __byte_perm(nextw, 0, 0x0123) : w1;
            wp += need ? 1 : 0;
            if (MODE == HUFF_FAST) {
                // predicated load instead of a branch (a lone warp cannot issue past an unresolved branch, and this
                // loop is bound by its latency per symbol): the destuffed data is followed by >= 16 zero bytes
                const uint32_t *ptr = src + min(wp, wmax + 2);
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\t@p ld.global.nc.u32 %0, [%1];\n\t}"
                             : "+r"(nextw) : "l"(ptr), "r"((int)need));
            } else
            if (need) nextw = wp < wmax ? src[wp] : 0;
            huff_prefetch(src, wp, wmax, need);
        }
        const bool is_dc = k == 0;
        const int t = (comp ? 2 : 0) + (is_dc ? 0 : 1);
        const uint32_t win = __funnelshift_l(w1, w0, bp);
        const uint32_t win16 = win >> 16;
        const uint32_t e = lds_u16(lut_base + (uint32_t)t * (kLutEntries * 2) + ((win >> (32 - kLook)) << 1));
        int len = (int)(e >> 8), sym = (int)(e & 255);
        if (e == 0) {                                    // code longer than kLook bits (rare)
            int l = kLook + 1;
#pragma unroll
            for (int j = 0; j < 16 - kLook - 1; j++) l += win16 >= T.limit[t][j] ? 1 : 0;
            const bool valid = win16 < T.limit[t][16 - kLook - 1];
            const int idx = (int)(win16 >> (16 - l)) + T.valoff[t][l - kLook - 1];
            sym = valid ? T.vals[t][idx & 255] : 0;                  // invalid code: libjpeg warns, returns 0
            len = l;
        }
        const int s = sym & 15, r = sym >> 4;
        const uint32_t bits = ((win << len) >> 1) >> (31 - s);
        bp += len + s;
        const int v = (s && bits < (1u << (s - 1))) ? (int)bits - (1 << s) + 1 : (int)bits;   // EXTEND
        // DC: difference to the predictor of the component; AC: run r, value v; (0,0) end of block; (15,0) 16 zeros
        pred += is_dc ? v : 0;
        const int pos = is_dc ? 0 : k + r;
        if (MODE == HUFF_FAST) {
            // no branch around the store: what is not to be kept goes to the stream's spare block
            int16_t *dst = (is_dc || (s != 0 && pos < 64)) ? blk + pos : trash;
            *dst = (int16_t)(is_dc ? pred : v);
        } else
        if (is_dc || (s != 0 && pos < 64)) blk[pos] = (int16_t)(is_dc ? pred : v);
        k = (is_dc || s != 0) ? pos + 1 : (r == 15 ? k + 16 : 64);
        const bool block_end = k >= 64;
        if (block_end) {
            k = 0;
            kept += keep ? 64 : 0;
            left--;
            const bool mcu_end = slot + 1 == nslots;
            slot = mcu_end ? 0 : slot + 1;
            const int ncomp = (int)(slot_info >> (4 * slot) & 3);
            if (ncomp != comp) {
                // the scan interleaves the components 0 -> 1 -> 2 -> 0 (parse_stream checks the order):
                // park the predictor of the component that ends, fetch the one that begins
                const int t0 = pred;
                pred = pred1; pred1 = pred2; pred2 = t0;
                comp = ncomp;
            }
            if (MODE == HUFF_FAST) {
                if (mcu_end && --run == 0) {
                    keep = !keep;
                    run = keep ? run_keep : run_skip;
                }
            } else
            if (mcu_end) {                                            // one MCU in every nslots blocks
                if (--run == 0) {
                    keep = !keep;
                    run = keep ? run_keep : run_skip;
                }
                // jdhuff.c insufficient_data: a symbol of this MCU needed bits past the end of the data (a
                // stream cut short).  libjpeg finishes that MCU on zero bits, as was just done, and leaves the
                // following MCUs untouched -- all zero, mid-grey after the IDCT -- until a restart re-arms it.
                bool dry = (uint32_t)(wp - 2) * 32u + (uint32_t)bp > avail_bits;
                if (restart && ++in_interval == restart) {            // jdhuff.c process_restart
                    dry = false;
                    in_interval = 0;
                    bp = (bp + 7) & ~7;                               // to the byte boundary
                    if (bp >= 64) {                                   // (the loop's refill covers bp < 64)
                        bp -= 32;
                        w0 = w1;
                        w1 = __byte_perm(nextw, 0, 0x0123);
                        wp++;
                        nextw = wp < wmax ? src[wp] : 0;
                    }
                    pred = pred1 = pred2 = 0;
                }
                while (dry && left > 0) {                             // skip one MCU (rare: damaged streams only)
                    kept += keep ? nslots * 64 : 0;
                    left -= nslots;
                    if (--run == 0) {
                        keep = !keep;
                        run = keep ? run_keep : run_skip;
                    }
                    if (restart && ++in_interval == restart) {
                        in_interval = 0;
                        dry = false;
                        pred = pred1 = pred2 = 0;
                    }
                }
            }
            blk = keep ? kept : trash;
        }
    }
    // a symbol needed bits past the end of the data (a stream cut short): the careful variant decides what libjpeg
    // would have left in the coefficients
    if (MODE == HUFF_FAST) flags[img] = (uint32_t)(wp - 2) * 32u + (uint32_t)bp > avail_bits ? 1u : 0u;
}

// ---------------------------------------------------------------------------------------------
// kernel 2: dequantisation + inverse DCT (libjpeg jidctint.c jpeg_idct_islow), one thread per block
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// One 8-point pass of the Loeffler-Ligtenberg-Moschytz flow graph with libjpeg's 13-bit constants.
__device__ __forceinline__ void idct8(int i0, int i1, int i2, int i3, int i4, int i5, int i6, int i7,
                                      int shift, int (&o)[8])
{
    int z1 = (i2 + i6) * 4433;
    int tmp2 = z1 + i6 * (-15137);
    int tmp3 = z1 + i2 * 6270;
    int tmp0 = (int)((unsigned)(i0 + i4) << 13);
    int tmp1 = (int)((unsigned)(i0 - i4) << 13);
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = i7; tmp1 = i5; tmp2 = i3; tmp3 = i1;
    z1 = tmp0 + tmp3;
    int z2 = tmp1 + tmp2, z3 = tmp0 + tmp2, z4 = tmp1 + tmp3;
    const int z5 = (z3 + z4) * 9633;
    tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    o[0] = descale(tmp10 + tmp3, shift); o[7] = descale(tmp10 - tmp3, shift);
    o[1] = descale(tmp11 + tmp2, shift); o[6] = descale(tmp11 - tmp2, shift);
    o[2] = descale(tmp12 + tmp1, shift); o[5] = descale(tmp12 - tmp1, shift);
    o[3] = descale(tmp13 + tmp0, shift); o[4] = descale(tmp13 - tmp0, shift);
}

constexpr int IDCT_THREADS = 128;

struct BlockRect { int bx0[2], bx1[2], by0[2], by1[2]; };   // [0] luma, [1] chroma: blocks to transform

__global__ void __launch_bounds__(IDCT_THREADS)
melf_jpeg_idct_kernel(const uint8_t *__restrict__ blob, const JpegDesc *__restrict__ desc,
                      const int16_t *__restrict__ coef, const JpegGeom g, const BlockRect rect,
                      const McuRect R, uint8_t *__restrict__ planes)
{
    constexpr uint8_t ZZ[64] = {
        0, 1, 5, 6, 14, 15, 27, 28, 2, 4, 7, 13, 16, 26, 29, 42, 3, 8, 12, 17, 25, 30, 41, 43,
        9, 11, 18, 24, 31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60,
        21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};
    const int img = blockIdx.y;
    // thread = one block of the ones the colour stage will read: the luma rectangle first, then the
    // chroma rectangle of each chroma component
    int t = blockIdx.x * IDCT_THREADS + threadIdx.x;
    const int rw0 = rect.bx1[0] - rect.bx0[0] + 1, n0 = rw0 * (rect.by1[0] - rect.by0[0] + 1);
    const int rw1 = rect.bx1[1] - rect.bx0[1] + 1, n1 = rw1 * (rect.by1[1] - rect.by0[1] + 1);
    if (t >= n0 + (g.ncomp == 3 ? 2 * n1 : 0)) return;
    int comp = 0;
    if (t >= n0) {
        t -= n0;
        comp = 1 + (t >= n1 ? 1 : 0);
        t -= t >= n1 ? n1 : 0;
    }
    const int rc = comp ? 1 : 0;
    const int rw = comp ? rw1 : rw0;
    const int by = rect.by0[rc] + t / rw, bx = rect.bx0[rc] + t % rw;
    const int bw = comp == 0 ? g.blocks_w[0] : g.blocks_w[1];
    const JpegDesc d = desc[img];
    const uint8_t *q = blob + d.base + (comp == 0 ? d.quant_off[0] : comp == 1 ? d.quant_off[1] : d.quant_off[2]);
    const bool q16 = d.quant16 >> comp & 1;

    // 64 coefficients (zig-zag order in memory) -> registers, dequantised
    // the Huffman stage stores blocks in scan order: MCU-major, component slots inside an MCU
    const int ch = comp ? 1 : g.hs, cv = comp ? 1 : g.vs;
    const int slot0 = comp == 0 ? 0 : g.hs * g.vs + comp - 1;
    const int scan_idx = ((by / cv - R.my0) * R.mw + (bx / ch - R.mx0)) * g.nslots + slot0 + (by % cv) * ch + bx % ch;
    const uint4 *src = reinterpret_cast<const uint4 *>(coef + ((size_t)img * R.stride_blocks + scan_idx) * 64);
    int z[64];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint4 v = __ldg(src + j);
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            z[8 * j + 2 * k] = (int)(int16_t)(w4[k] & 0xFFFF);
            z[8 * j + 2 * k + 1] = (int)(int16_t)(w4[k] >> 16);
        }
    }
#pragma unroll
    for (int j = 0; j < 64; j++) {
        const int qv = q16 ? (q[2 * j] << 8 | q[2 * j + 1]) : q[j];
        z[j] *= qv;
    }
    // pass 1: columns (natural index = 8 * row + col)
    int ws[64];
#pragma unroll
    for (int c = 0; c < 8; c++) {
        int o[8];
        idct8(z[ZZ[c]], z[ZZ[8 + c]], z[ZZ[16 + c]], z[ZZ[24 + c]], z[ZZ[32 + c]], z[ZZ[40 + c]],
              z[ZZ[48 + c]], z[ZZ[56 + c]], 13 - 2, o);
#pragma unroll
        for (int r = 0; r < 8; r++) ws[8 * r + c] = o[r];
    }
    // pass 2: rows, range limit (+128, clamp), 8 samples = one 8-byte store
    const int stride = bw * 8;
    uint8_t *dst = planes + (size_t)img * g.plane_bytes +
                   (comp == 0 ? g.plane_off[0] : comp == 1 ? g.plane_off[1] : g.plane_off[2]) +
                   (size_t)by * 8 * stride + bx * 8;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        int o[8];
        idct8(ws[8 * r], ws[8 * r + 1], ws[8 * r + 2], ws[8 * r + 3], ws[8 * r + 4], ws[8 * r + 5],
              ws[8 * r + 6], ws[8 * r + 7], 13 + 2 + 3, o);
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            lo |= (uint32_t)min(max(o[c] + 128, 0), 255) << (8 * c);
            hi |= (uint32_t)min(max(o[4 + c] + 128, 0), 255) << (8 * c);
        }
        *reinterpret_cast<uint2 *>(dst + (size_t)r * stride) = make_uint2(lo, hi);
    }
}

// ---------------------------------------------------------------------------------------------
// kernel 3: chroma upsampling (libjpeg jdsample.c, edge rows per jdmainct.c) + YCbCr -> BGR
// (jdcolor.c), 4 pixels of one row per thread, for the rectangle [x0, x0+cw) x [y0, y0+ch)
// ---------------------------------------------------------------------------------------------
struct ColorArgs {
    const uint8_t *planes;
    uint8_t *out;
    size_t row_stride, frame_stride;
    int x0, y0, cw, ch, n;
};

__device__ __forceinline__ int upsampled(const uint8_t *__restrict__ p, int stride, const JpegGeom &g,
                                         int x, int y)
{
    if (g.hs == 1) return p[(size_t)y * stride + x];
    const int cx = x >> 1, last = g.ds_w - 1;
    if (g.ds_w <= 2) return p[(size_t)(g.vs == 2 ? y >> 1 : y) * stride + cx];   // jinit_upsampler: replicate
    if (g.vs == 1) {                                                           // h2v1_fancy_upsample
        const uint8_t *r = p + (size_t)y * stride;
        const int v = r[cx];
        if (x & 1) return cx == last ? v : (3 * v + r[cx + 1] + 2) >> 2;
        return cx == 0 ? v : (3 * v + r[cx - 1] + 1) >> 2;
    }
    const int cy = y >> 1;                                                     // h2v2_fancy_upsample
    const int oy = min(max((y & 1) ? cy + 1 : cy - 1, 0), g.ds_h - 1);
    const uint8_t *r0 = p + (size_t)cy * stride, *r1 = p + (size_t)oy * stride;
    const int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) {
        if (cx == last) return (4 * cur + 7) >> 4;
        return (3 * cur + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    }
    if (cx == 0) return (4 * cur + 8) >> 4;
    return (3 * cur + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}

constexpr int COLOR_THREADS = 128;

__device__ __forceinline__ void ycc_to_bgr(int yy, int cb, int cr, uint8_t *px)
{
    cb -= 128;
    cr -= 128;
    const int r = yy + ((91881 * cr + 32768) >> 16);                       // FIX(1.40200)
    const int gg = yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16);         // FIX(0.34414), FIX(0.71414)
    const int b = yy + ((116130 * cb + 32768) >> 16);                      // FIX(1.77200)
    px[0] = (uint8_t)min(max(b, 0), 255);
    px[1] = (uint8_t)min(max(gg, 0), 255);
    px[2] = (uint8_t)min(max(r, 0), 255);
}

// h2v2 row pair -> the 4 upsampled samples of pixels 4T .. 4T+3 (chroma columns 2T, 2T+1 and their
// neighbours 2T-1, 2T+2; the first and last column rules of h2v2_fancy_upsample at the edges)
__device__ __forceinline__ void h2v2_four(const uint8_t *__restrict__ r0, const uint8_t *__restrict__ r1,
                                          int c0, int last, int (&o)[4])
{
    const uint32_t a = *reinterpret_cast<const uint16_t *>(r0 + c0);       // c0 is even: aligned pair
    const uint32_t b = *reinterpret_cast<const uint16_t *>(r1 + c0);
    const int s0 = 3 * (int)(a & 255) + (int)(b & 255), s1 = 3 * (int)(a >> 8) + (int)(b >> 8);
    const int sl = c0 > 0 ? 3 * r0[c0 - 1] + r1[c0 - 1] : 0;
    const int sr = c0 + 2 <= last ? 3 * r0[c0 + 2] + r1[c0 + 2] : 0;
    o[0] = c0 == 0 ? (4 * s0 + 8) >> 4 : (3 * s0 + sl + 8) >> 4;
    o[1] = c0 == last ? (4 * s0 + 7) >> 4 : (3 * s0 + s1 + 7) >> 4;
    o[2] = (3 * s1 + s0 + 8) >> 4;
    o[3] = c0 + 1 == last ? (4 * s1 + 7) >> 4 : (3 * s1 + sr + 7) >> 4;
}

__global__ void __launch_bounds__(COLOR_THREADS)
melf_jpeg_color_kernel(const ColorArgs a, const JpegGeom g)
{
    // thread = 4 pixels of one row at an image column that is a multiple of 4
    const int t0 = a.x0 >> 2;
    const int T = t0 + blockIdx.x * blockDim.x + threadIdx.x;
    const int y = a.y0 + blockIdx.y, img = blockIdx.z;
    const int X = 4 * T;
    if (X >= a.x0 + a.cw) return;
    const uint8_t *pl = a.planes + (size_t)img * g.plane_bytes;
    const int ystride = g.blocks_w[0] * 8;
    const uint8_t *yp = pl + g.plane_off[0] + (size_t)y * ystride;
    const int cstride = g.blocks_w[1] * 8;
    uint8_t px[12];
    const uint32_t y4 = *reinterpret_cast<const uint32_t *>(yp + X);       // the plane is padded to whole blocks
    if (g.ncomp == 3 && g.hs == 2 && g.vs == 2 && g.ds_w > 2 && X + 3 <= 2 * g.ds_w - 1) {
        const int cy = y >> 1;
        const int oy = min(max((y & 1) ? cy + 1 : cy - 1, 0), g.ds_h - 1);
        int cb[4], cr[4];
        const uint8_t *pb = pl + g.plane_off[1], *pr = pl + g.plane_off[2];
        h2v2_four(pb + (size_t)cy * cstride, pb + (size_t)oy * cstride, X >> 1, g.ds_w - 1, cb);
        h2v2_four(pr + (size_t)cy * cstride, pr + (size_t)oy * cstride, X >> 1, g.ds_w - 1, cr);
#pragma unroll
        for (int k = 0; k < 4; k++) ycc_to_bgr((int)(y4 >> (8 * k) & 255), cb[k], cr[k], px + 3 * k);
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = min(X + k, g.w - 1);
            const int yy = yp[x];
            if (g.ncomp == 3)
                ycc_to_bgr(yy, upsampled(pl + g.plane_off[1], cstride, g, x, y),
                           upsampled(pl + g.plane_off[2], cstride, g, x, y), px + 3 * k);
            else
                px[3 * k] = px[3 * k + 1] = px[3 * k + 2] = (uint8_t)yy;
        }
    }
    uint8_t *o = a.out + (size_t)img * a.frame_stride + (size_t)blockIdx.y * a.row_stride;
    const int rel = X - a.x0;                       // column of the first pixel inside the rectangle
    if (rel >= 0 && rel + 4 <= a.cw && ((reinterpret_cast<uintptr_t>(o + 3 * rel) & 3) == 0)) {
        uint32_t *o32 = reinterpret_cast<uint32_t *>(o + 3 * rel);
#pragma unroll
        for (int k = 0; k < 3; k++)
            o32[k] = px[4 * k] | px[4 * k + 1] << 8 | px[4 * k + 2] << 16 | (uint32_t)px[4 * k + 3] << 24;
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (rel + k >= 0 && rel + k < a.cw) {
                o[3 * (rel + k)] = px[3 * k];
                o[3 * (rel + k) + 1] = px[3 * k + 1];
                o[3 * (rel + k) + 2] = px[3 * k + 2];
            }
    }
}

template <typename T>
cudaError_t grow(T **p, size_t *cap, size_t need)
{
    if (need <= *cap) return cudaSuccess;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    const size_t bytes = need + need / 8;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(p), bytes);
    if (e == cudaSuccess) *cap = bytes;
    return e;
}

}  // namespace

struct MelfJpegState {
    uint8_t *d_blob = nullptr;      size_t blob_cap = 0;
    uint8_t *d_clean = nullptr;     size_t clean_cap = 0;
    JpegDesc *d_desc = nullptr;     size_t desc_cap = 0;
    uint32_t *d_clean_len = nullptr; size_t len_cap = 0;
    int16_t *d_coef = nullptr;      size_t coef_cap = 0;
    uint8_t *d_planes = nullptr;    size_t planes_cap = 0;
    uint8_t *h_stage = nullptr;     size_t stage_cap = 0;     // pinned: descriptors + streams of one call
    cudaEvent_t uploaded = nullptr;
    bool upload_pending = false;
};

void melf_jpeg_state_free(MelfJpegState *s)
{
    if (!s) return;
    cudaFree(s->d_blob); cudaFree(s->d_clean); cudaFree(s->d_desc); cudaFree(s->d_clean_len);
    cudaFree(s->d_coef); cudaFree(s->d_planes);
    if (s->h_stage) cudaFreeHost(s->h_stage);
    if (s->uploaded) cudaEventDestroy(s->uploaded);
    delete s;
}

int melf_jpeg_probe(const uint8_t *jpeg, size_t len, int *h, int *w, int *ncomp, int *hs, int *vs,
                    std::string *err)
{
    JpegDesc d{};
    JpegGeom g{};
    const char *e = parse_stream(jpeg, len, &d, &g, true);
    if (e) {
        if (err) *err = e;
        return std::strncmp(e, "unsupported", 11) == 0 ? 3 : 1;
    }
    if (h) *h = g.h;
    if (w) *w = g.w;
    if (ncomp) *ncomp = g.ncomp;
    if (hs) *hs = g.hs;
    if (vs) *vs = g.vs;
    return 0;
}

#define JPEG_CUDA(call)                                                            \
    do {                                                                           \
        cudaError_t e_ = (call);                                                   \
        if (e_ != cudaSuccess) {                                                   \
            if (err) *err = std::string(#call) + ": " + cudaGetErrorString(e_);    \
            return 2;                                                              \
        }                                                                          \
    } while (0)

int melf_jpeg_decode_range(MelfJpegState **state, const MelfJpegBatch &batch, int first, int count,
                           int h, int w, int x0, int y0, int cw, int ch, uint8_t *d_out,
                           size_t row_stride, size_t frame_stride, cudaStream_t stream,
                           std::string *err, uint64_t *launches, cudaEvent_t (*spans)[2], cudaStream_t copy_stream)
{
    if (count <= 0) return 0;
    if (!*state) *state = new MelfJpegState();
    MelfJpegState &S = **state;
    if (!S.uploaded) JPEG_CUDA(cudaEventCreateWithFlags(&S.uploaded, cudaEventDisableTiming));

    // ---- host: marker walk, descriptors, staging copy ----
    const uint64_t blob0 = batch.offsets[first];
    const size_t blob_bytes = (size_t)(batch.offsets[first + count] - blob0);
    // descriptors, then the lane order of the Huffman stage (u32 per stream), then the streams
    const size_t order_off = ((size_t)count * sizeof(JpegDesc) + 15) & ~(size_t)15;
    const size_t desc_bytes = (order_off + (size_t)count * 4 + 255) & ~(size_t)255;
    if (S.upload_pending) {             // the staging buffer is still the source of the previous upload
        JPEG_CUDA(cudaEventSynchronize(S.uploaded));
        S.upload_pending = false;
    }
    // a caller blob that is already page-locked is uploaded from where it lies; otherwise it is
    // gathered into the pinned staging buffer behind the descriptors
    cudaPointerAttributes pattr{};
    const bool direct = cudaPointerGetAttributes(&pattr, batch.jpeg) == cudaSuccess &&
                        pattr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    const size_t stage_need = desc_bytes + (direct ? 0 : blob_bytes) + 16;
    if (stage_need > S.stage_cap) {
        if (S.h_stage) cudaFreeHost(S.h_stage);
        S.h_stage = nullptr;
        S.stage_cap = 0;
        const size_t bytes = stage_need * 5 / 4;
        JPEG_CUDA(cudaMallocHost(reinterpret_cast<void **>(&S.h_stage), bytes));
        S.stage_cap = bytes;
    }
    JpegDesc *descs = reinterpret_cast<JpegDesc *>(S.h_stage);
    JpegGeom g{};
    size_t clean_bytes = 0;
    bool shared_tables = true;
    // marker walk of streams [i0, i1): descriptors, DHT comparison against stream 0 (whose descriptor and the
    // geometry `g` exist by then); reports the first stream that fails
    struct WalkResult { int bad = -1; int code = 0; std::string text; bool shared = true; };
    auto walk = [&](int i0, int i1, WalkResult *r) {
        for (int i = i0; i < i1; i++) {
            const uint64_t o0 = batch.offsets[first + i], o1 = batch.offsets[first + i + 1];
            const char *e = nullptr;
            JpegDesc &d = descs[i];
            if (o1 < o0 || o1 - o0 > 0xFFFFFFF0ull) {
                e = "bad stream offsets";
            } else {
                std::memset(&d, 0, sizeof d);
                e = parse_stream(batch.jpeg + o0, (size_t)(o1 - o0), &d, &g, i == 0);
            }
            if (e) {
                r->bad = i;
                r->code = std::strncmp(e, "unsupported", 11) == 0 ? 3 : 1;
                r->text = e;
                return;
            }
            if (r->shared && i > 0) {
                const uint8_t *a = batch.jpeg + batch.offsets[first], *b = batch.jpeg + o0;
                for (int t = 0; t < 4 && r->shared; t++) {
                    const uint8_t *ta = a + descs[0].huff_off[t], *tb = b + d.huff_off[t];
                    int total = 0;
                    for (int k = 0; k < 16; k++) total += ta[k];
                    r->shared = std::memcmp(ta, tb, 16 + (size_t)total) == 0;
                }
            }
            d.base = o0 - blob0;
        }
    };
    // stream 0 first (it defines the geometry), the rest on a few host threads when the chunk is large
    static const int walk_threads = [] {
        const char *e = std::getenv("MELF_JPEG_HOST_THREADS");
        int v = e ? std::atoi(e) : (int)std::thread::hardware_concurrency() / 2;
        // one process per GPU on the same host (torchrun): share the cores
        const char *lws = std::getenv("LOCAL_WORLD_SIZE");
        if (!e && lws && std::atoi(lws) > 1) v /= std::atoi(lws);
        return v < 1 ? 1 : v > 8 ? 8 : v;
    }();
    const int nt = count >= 2048 ? walk_threads : 1;
    std::vector<WalkResult> results((size_t)nt + 1);
    walk(0, 1, &results[0]);
    if (results[0].bad < 0) {
        if (nt == 1) {
            walk(1, count, &results[1]);
        } else {
            std::vector<std::thread> pool;
            const int per = (count - 1 + nt - 1) / nt;
            for (int t = 0; t < nt; t++) {
                const int i0 = 1 + t * per, i1 = i0 + per < count ? i0 + per : count;
                if (i0 < i1) pool.emplace_back(walk, i0, i1, &results[(size_t)t + 1]);
            }
            for (std::thread &th : pool) th.join();
        }
    }
    for (const WalkResult &r : results) {                 // ranges are in stream order: the first failure wins
        if (r.bad >= 0) {
            if (err) *err = r.text == "bad stream offsets" ? r.text : "stream " + std::to_string(first + r.bad) + ": " + r.text;
            return r.code;
        }
        shared_tables = shared_tables && r.shared;
    }
    for (int i = 0; i < count; i++) {
        descs[i].clean_off = clean_bytes;
        clean_bytes += (((size_t)descs[i].scan_len + 15) & ~(size_t)15) + 32;
    }
    // Huffman stage: one lane per stream, a warp runs as long as its longest stream.  The lanes take the
    // streams in order of entropy-coded length (the host has it from the marker walk), so the streams of a
    // warp end together -- in stream order ncu counted 10.6 live lanes of 16 (profiles/r01_v19_ncu_jpeg_huffman.txt).
    // Nothing else moves: coefficients, planes and frames stay indexed by stream.
    {
        uint32_t *order = reinterpret_cast<uint32_t *>(S.h_stage + order_off);
        for (int i = 0; i < count; i++) order[i] = (uint32_t)i;
        // ... in buckets of 1 KB (MELF_JPEG_SORT=<shift>, default 10; 0 = stream order, 1 = exact lengths): inside a
        // bucket the streams keep their order.  A finer order buys nothing (lanes within 3 % of each other), and an
        // exact one puts byte-identical streams -- a benchmark's repeated files -- into one warp, whose lanes then
        // hit the same L2 slices with every store (profiles/r02_huffman_variants.txt: slower than stream order).
        static const int sort_shift = std::getenv("MELF_JPEG_SORT") ? std::atoi(std::getenv("MELF_JPEG_SORT")) : 10;
        if (sort_shift > 0) {
            const int sh = sort_shift == 1 ? 0 : sort_shift;
            std::stable_sort(order, order + count, [&](uint32_t a, uint32_t b) {
                return (descs[a].scan_len >> sh) > (descs[b].scan_len >> sh); });
        }
    }
    if (g.w != w || g.h != h) {
        if (err) *err = "the streams decode to " + std::to_string(g.w) + "x" + std::to_string(g.h) +
                        " pixels, not the " + std::to_string(w) + "x" + std::to_string(h) + " asked for";
        return 1;
    }
    if (x0 < 0 || y0 < 0 || cw < 1 || ch < 1 || x0 + cw > w || y0 + ch > h || row_stride < (size_t)cw * 3) {
        if (err) *err = "bad output rectangle";
        return 1;
    }
    if (!direct) std::memcpy(S.h_stage + desc_bytes, batch.jpeg + blob0, blob_bytes);

    // ---- blocks under the output rectangle, and the MCUs that hold them ----
    BlockRect br;
    br.bx0[0] = x0 >> 3; br.bx1[0] = (x0 + cw - 1) >> 3;
    br.by0[0] = y0 >> 3; br.by1[0] = (y0 + ch - 1) >> 3;
    {   // chroma samples under the rectangle plus the one-sample apron of the triangle filter
        const int cx0 = x0 / g.hs - 1, cx1 = (x0 + cw - 1) / g.hs + 1, cy0 = y0 / g.vs - 1, cy1 = (y0 + ch - 1) / g.vs + 1;
        br.bx0[1] = (cx0 < 0 ? 0 : cx0) >> 3; br.bx1[1] = cx1 >> 3;
        br.by0[1] = (cy0 < 0 ? 0 : cy0) >> 3; br.by1[1] = cy1 >> 3;
        if (br.bx1[1] > g.blocks_w[g.ncomp == 3 ? 1 : 0] - 1) br.bx1[1] = g.blocks_w[g.ncomp == 3 ? 1 : 0] - 1;
        if (br.by1[1] > g.blocks_h[g.ncomp == 3 ? 1 : 0] - 1) br.by1[1] = g.blocks_h[g.ncomp == 3 ? 1 : 0] - 1;
    }
    McuRect mr;
    {
        const bool chroma = g.ncomp == 3;
        mr.mx0 = br.bx0[0] / g.hs; mr.mx1 = br.bx1[0] / g.hs;
        mr.my0 = br.by0[0] / g.vs; mr.my1 = br.by1[0] / g.vs;
        if (chroma) {
            if (br.bx0[1] < mr.mx0) mr.mx0 = br.bx0[1];
            if (br.bx1[1] > mr.mx1) mr.mx1 = br.bx1[1];
            if (br.by0[1] < mr.my0) mr.my0 = br.by0[1];
            if (br.by1[1] > mr.my1) mr.my1 = br.by1[1];
        }
        mr.mw = mr.mx1 - mr.mx0 + 1;
        mr.stride_blocks = mr.mw * (mr.my1 - mr.my0 + 1) * g.nslots + 1;
    }

    // ---- device buffers ----
    const size_t coef_elems = (size_t)count * mr.stride_blocks * 64;
    JPEG_CUDA(grow(&S.d_blob, &S.blob_cap, desc_bytes + blob_bytes + 16));
    JPEG_CUDA(grow(&S.d_clean, &S.clean_cap, clean_bytes + 64));
    JPEG_CUDA(grow(&S.d_clean_len, &S.len_cap, (size_t)count * 8));     // lengths, then the redo flags of the Huffman stage
    JPEG_CUDA(grow(&S.d_coef, &S.coef_cap, coef_elems * 2));
    JPEG_CUDA(grow(&S.d_planes, &S.planes_cap, (size_t)count * g.plane_bytes));
    (void)S.d_desc; (void)S.desc_cap;

    // The upload goes on the caller's copy stream when there is one: the chunks of a batch then cross PCIe back to
    // back in issue order, whatever the chunk streams are busy with (their copies used to start late behind other
    // chunks' kernels: profiles/r02_ingest_trace_65536.txt), and `stream` waits for the event.  The caller makes
    // sure the previous work on this state has finished before it calls again (it reads d_blob).
    const cudaStream_t cs = copy_stream ? copy_stream : stream;
    if (spans) cudaEventRecord(spans[0][0], cs);
    if (direct) {
        JPEG_CUDA(cudaMemcpyAsync(S.d_blob, S.h_stage, desc_bytes, cudaMemcpyHostToDevice, cs));
        JPEG_CUDA(cudaMemcpyAsync(S.d_blob + desc_bytes, batch.jpeg + blob0, blob_bytes, cudaMemcpyHostToDevice, cs));
    } else {
        JPEG_CUDA(cudaMemcpyAsync(S.d_blob, S.h_stage, desc_bytes + blob_bytes, cudaMemcpyHostToDevice, cs));
    }
    JPEG_CUDA(cudaEventRecord(S.uploaded, cs));
    S.upload_pending = true;
    if (spans) cudaEventRecord(spans[0][1], cs);
    const JpegDesc *d_desc = reinterpret_cast<const JpegDesc *>(S.d_blob);
    const uint32_t *d_order = reinterpret_cast<const uint32_t *>(S.d_blob + order_off);
    const uint8_t *d_streams = S.d_blob + desc_bytes;
    JPEG_CUDA(cudaMemsetAsync(S.d_coef, 0, coef_elems * 2, stream));        // (under the upload when that has its own stream)
    if (cs != stream) JPEG_CUDA(cudaStreamWaitEvent(stream, S.uploaded, 0));

    if (spans) cudaEventRecord(spans[1][0], stream);
    melf_jpeg_destuff_kernel<<<count, DESTUFF_THREADS, 0, stream>>>(d_streams, d_desc, S.d_clean, S.d_clean_len);
    if (spans) { cudaEventRecord(spans[1][1], stream); cudaEventRecord(spans[2][0], stream); }
    // lanes per warp: few enough that the batch still reaches every SM sub-partition
    int G = 16;                                  // measured: profiles/r01_v11_jpeg_sweep.txt
    static const int min_warps = std::getenv("MELF_JPEG_MIN_WARPS") ? std::atoi(std::getenv("MELF_JPEG_MIN_WARPS")) : 256;
    while (G > 2 && count / G < min_warps) G >>= 1;
    static const int forced = std::getenv("MELF_JPEG_LANES") ? std::atoi(std::getenv("MELF_JPEG_LANES")) : 0;
    if (forced == 1 || forced == 2 || forced == 4 || forced == 8 || forced == 16 || forced == 32) G = forced;
    const int ctas = (count + G - 1) / G;
    const size_t hsmem = sizeof(HuffSmem) * (shared_tables ? 1 : G);
    // chunks without restart intervals take the lean loop + the redo pass for streams that ran out of data
    // (MELF_JPEG_FAST=0: the careful loop for everything, A/B runs)
    bool any_restart = false;
    for (int i = 0; i < count; i++) any_restart = any_restart || descs[i].restart != 0;
    static const bool fast_allowed = !(std::getenv("MELF_JPEG_FAST") && !std::atoi(std::getenv("MELF_JPEG_FAST")));
    const bool fast = fast_allowed && !any_restart;
    uint32_t *d_flags = S.d_clean_len + count;
#define MELF_HUFF_GO(GG, SH, MODE)                                                                        \
    melf_jpeg_huffman_kernel<GG, SH, MODE><<<ctas, 32, hsmem, stream>>>(                                  \
        d_streams, S.d_clean, d_desc, d_order, S.d_clean_len, d_flags, count, g, mr, S.d_coef)
    static const int cta_warps = std::getenv("MELF_JPEG_CTA_WARPS") ? std::atoi(std::getenv("MELF_JPEG_CTA_WARPS")) : 8;   // 1, 4, 8, 16 measured: profiles/r02_huffman_variants.txt
#define MELF_HUFF_GO4(GG)                                                                                 \
    if (cta_warps == 16)                                                                                  \
        melf_jpeg_huffman_kernel<GG, true, HUFF_FAST, 16><<<(ctas + 15) / 16, 512, hsmem, stream>>>(     \
            d_streams, S.d_clean, d_desc, d_order, S.d_clean_len, d_flags, count, g, mr, S.d_coef);       \
    else if (cta_warps == 8)                                                                              \
        melf_jpeg_huffman_kernel<GG, true, HUFF_FAST, 8><<<(ctas + 7) / 8, 256, hsmem, stream>>>(        \
            d_streams, S.d_clean, d_desc, d_order, S.d_clean_len, d_flags, count, g, mr, S.d_coef);       \
    else                                                                                                  \
        melf_jpeg_huffman_kernel<GG, true, HUFF_FAST, 4><<<(ctas + 3) / 4, 128, hsmem, stream>>>(        \
            d_streams, S.d_clean, d_desc, d_order, S.d_clean_len, d_flags, count, g, mr, S.d_coef)
#define MELF_HUFF_LAUNCH(GG)                                                                              \
    case GG:                                                                                              \
        if (shared_tables) {                                                                              \
            if (fast && cta_warps >= 4 && !(MELF_HUFF_DEBUG & 4)) { MELF_HUFF_GO4(GG); MELF_HUFF_GO(GG, true, HUFF_REDO); } else \
            if (fast) { MELF_HUFF_GO(GG, true, HUFF_FAST); if (!(MELF_HUFF_DEBUG & 4)) MELF_HUFF_GO(GG, true, HUFF_REDO); }           \
            else MELF_HUFF_GO(GG, true, HUFF_CAREFUL);                                                    \
        } else {                                                                                          \
            JPEG_CUDA(cudaFuncSetAttribute(melf_jpeg_huffman_kernel<GG, false, HUFF_CAREFUL>,             \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));     \
            JPEG_CUDA(cudaFuncSetAttribute(melf_jpeg_huffman_kernel<GG, false, HUFF_FAST>,                \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));     \
            JPEG_CUDA(cudaFuncSetAttribute(melf_jpeg_huffman_kernel<GG, false, HUFF_REDO>,                \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));     \
            if (fast) { MELF_HUFF_GO(GG, false, HUFF_FAST); MELF_HUFF_GO(GG, false, HUFF_REDO); }         \
            else MELF_HUFF_GO(GG, false, HUFF_CAREFUL);                                                   \
        }                                                                                                 \
        break;
    switch (G) {
        MELF_HUFF_LAUNCH(32)
        MELF_HUFF_LAUNCH(16)
        MELF_HUFF_LAUNCH(8)
        MELF_HUFF_LAUNCH(4)
        MELF_HUFF_LAUNCH(2)
        MELF_HUFF_LAUNCH(1)
    }
#undef MELF_HUFF_GO
#undef MELF_HUFF_GO4
#if MELF_HUFF_DEBUG & 4
    if (const char *dump = std::getenv("MELF_HUFF_DUMP")) {
        std::vector<uint32_t> hf((size_t)count), ho((size_t)count);
        cudaStreamSynchronize(stream);
        cudaMemcpy(hf.data(), d_flags, (size_t)count * 4, cudaMemcpyDeviceToHost);
        if (FILE *fp = std::fopen(dump, "w")) {
            const uint32_t *order = reinterpret_cast<const uint32_t *>(S.h_stage + order_off);
            for (int i = 0; i < count; i++)
                std::fprintf(fp, "%d %u %u %u %u\n", i, order[i], descs[order[i]].scan_len, (hf[order[i]] >> 16) << 9, (hf[order[i]] & 0xffffu) << 2);
            std::fclose(fp);
        }
    }
#endif
#undef MELF_HUFF_LAUNCH
    if (spans) { cudaEventRecord(spans[2][1], stream); cudaEventRecord(spans[3][0], stream); }
    const int idct_blocks = (br.bx1[0] - br.bx0[0] + 1) * (br.by1[0] - br.by0[0] + 1) +
                            (g.ncomp == 3 ? 2 * (br.bx1[1] - br.bx0[1] + 1) * (br.by1[1] - br.by0[1] + 1) : 0);
    dim3 igrid((idct_blocks + IDCT_THREADS - 1) / IDCT_THREADS, count);
    melf_jpeg_idct_kernel<<<igrid, IDCT_THREADS, 0, stream>>>(d_streams, d_desc, S.d_coef, g, br, mr, S.d_planes);
    if (spans) { cudaEventRecord(spans[3][1], stream); cudaEventRecord(spans[4][0], stream); }
    ColorArgs ca{S.d_planes, d_out, row_stride, frame_stride, x0, y0, cw, ch, count};
    const int cgroups = ((x0 + cw + 3) >> 2) - (x0 >> 2);
    // narrow rectangles: 64-thread CTAs, so that no warp is launched just to return
    const int cthreads = cgroups <= 64 ? 64 : COLOR_THREADS;
    dim3 cgrid((cgroups + cthreads - 1) / cthreads, ch, count);
    melf_jpeg_color_kernel<<<cgrid, cthreads, 0, stream>>>(ca, g);
    if (spans) cudaEventRecord(spans[4][1], stream);
    JPEG_CUDA(cudaGetLastError());
    if (launches) *launches += fast ? 5 : 4;
    return 0;
}
