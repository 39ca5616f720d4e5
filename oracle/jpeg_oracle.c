/* CPU oracle for the ingest row: baseline JPEG -> BGR exactly as
 * cv2.imread(filename) returns it (reference call site: meterelf/_image.py:49).
 *
 * TEST INFRASTRUCTURE ONLY -- used by tests/, __graft_entry__.smoke() and bench.py's CPU legs,
 * never by the product package.
 *
 * The algorithm lives in a third-party dependency that is absent from /root/reference:
 * OpenCV's bundled libjpeg-turbo (cv2.getBuildInformation(): "build-libjpeg-turbo (ver 3.1.2-70)"),
 * driven by OpenCV's JpegDecoder with libjpeg defaults: dct_method = JDCT_ISLOW,
 * do_fancy_upsampling = TRUE, out_color_space = JCS_EXT_BGR.  This file restates the published
 * algorithm of that decode path:
 *   - sequential Huffman entropy decoding (ITU-T T.81 Annex F.2; libjpeg jdhuff.c),
 *   - dequantisation + the "slow but accurate" integer inverse DCT (Loeffler-Ligtenberg-Moschytz,
 *     13-bit constants, 2 bits of pass-1 precision; libjpeg jidctint.c),
 *   - "fancy" triangle-filter chroma upsampling for h2v1 and h2v2 with the edge rules of
 *     libjpeg jdsample.c / jdmainct.c (context rows above the first / below the last real row
 *     are copies of that row; first/last column special cases),
 *   - the 16-bit fixed point YCbCr -> RGB tables of libjpeg jdcolor.c.
 *   - files cut short: jdhuff.c insufficient_data (the MCU in which the data runs dry is finished on zero
 *     bits, the following MCUs stay all zero until a restart marker re-arms the decoder); pinned against
 *     cv2.imread on 89 truncated files (tests/golden/make_golden_jpeg_truncated.py).  Not restated:
 *     jpeg_resync_to_restart on the fake EOI of libjpeg's file source (truncated streams WITH restarts).
 * Pinned against cv2.imread on all 304 sample JPEGs of the reference (tests/golden/
 * make_golden_jpeg.py, run in the build container) -- see tests/test_jpeg_oracle.py.
 *
 * Supported: SOF0/SOF1 8-bit Huffman, one scan; 3 components with luma sampling 1x1, 2x1 or 2x2
 * and 1x1 chroma, or 1 component (grey, replicated into B, G, R); restart intervals.
 * Everything else returns a negative code (the product returns an error for the same inputs).
 * 3-component files that libjpeg reads as RGB (Adobe APP14 transform 0 without JFIF, or component
 * ids 'R','G','B') are decoded without the YCbCr conversion here; the product sends those files
 * through cv2.imread (melf_jpeg_info: "unsupported: RGB colour space").
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define JO_OK 0
#define JO_ERR_FORMAT (-1)       /* not a JPEG / truncated headers */
#define JO_ERR_UNSUPPORTED (-2)  /* progressive, arithmetic, 12-bit, exotic sampling, multi-scan */
#define JO_ERR_SPACE (-3)

static const uint8_t ZIGZAG[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
    41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

typedef struct {
    int present;
    uint16_t look[65536]; /* (length << 8) | symbol for the code at the top of a 16-bit window; 0 = invalid */
} HuffTable;

typedef struct {
    int id, hs, vs, tq, td, ta;
    int blocks_w, blocks_h;          /* padded to whole MCUs */
    int ds_w, ds_h;                  /* downsampled_width / downsampled_height */
    int16_t *coef;                   /* blocks_h * blocks_w * 64, natural order */
    uint8_t *plane;                  /* (blocks_h*8) x (blocks_w*8) */
} Component;

typedef struct {
    int w, h, ncomp, max_h, max_v, mcus_x, mcus_y, restart_interval;
    int rgb;                         /* libjpeg chose JCS_RGB: no YCbCr conversion (default_decompress_parms) */
    uint16_t quant[4][64];           /* natural order */
    int quant_present[4];
    HuffTable *dc[4], *ac[4];
    Component comp[3];
    const uint8_t *scan;             /* entropy-coded data */
    size_t scan_len;
} Decoder;

/* ---- bit reader (jdhuff.c: jpeg_fill_bit_buffer; zero bits after a marker) ---- */
typedef struct {
    const uint8_t *p, *end;
    uint64_t acc;
    int bits;
    int hit_marker;
    int real;          /* how many of `bits` are data; the rest is the zero padding after the data */
    int insufficient;  /* jdhuff.c insufficient_data: a symbol needed more bits than the data holds */
} BitReader;

static void br_fill(BitReader *br) {
    while (br->bits <= 48) {
        unsigned c = 0;
        if (!br->hit_marker && br->p < br->end) {
            c = *br->p;
            if (c == 0xFF) {
                if (br->p + 1 < br->end && br->p[1] == 0x00) {
                    br->p += 2;
                } else {
                    br->hit_marker = 1; /* leave p at the marker */
                    c = 0;
                }
            } else {
                br->p += 1;
            }
        } else {
            br->hit_marker = 1;
        }
        if (!br->hit_marker) br->real = br->bits + 8;      /* data bytes always sit in front of the padding */
        br->acc |= (uint64_t)c << (56 - br->bits);
        br->bits += 8;
    }
}

static inline unsigned br_peek16(BitReader *br) {
    if (br->bits < 16) br_fill(br);
    return (unsigned)(br->acc >> 48);
}

static inline void br_skip(BitReader *br, int n) {
    br->acc <<= n;
    br->bits -= n;
    if (n > br->real) {                 /* jpeg_fill_bit_buffer: nbits > bits_left and no more bytes */
        br->insufficient = 1;
        br->real = 0;
    } else {
        br->real -= n;
    }
}

static inline int br_get(BitReader *br, int n) {
    if (n == 0) return 0;
    if (br->bits < n) br_fill(br);
    int v = (int)(br->acc >> (64 - n));
    br_skip(br, n);
    return v;
}

/* T.81 F.2.2.1 EXTEND */
static inline int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

static int decode_symbol(BitReader *br, const HuffTable *t) {
    unsigned e = t->look[br_peek16(br)];
    if (!e) { br_skip(br, 16 < br->bits ? 16 : br->bits); return 0; } /* corrupt: libjpeg warns, returns 0 */
    br_skip(br, (int)(e >> 8));
    return (int)(e & 0xFF);
}

/* ---- headers (T.81 Annex B) ---- */
static int build_huff(HuffTable *t, const uint8_t *counts, const uint8_t *symbols, int nsym) {
    memset(t->look, 0, sizeof t->look);
    unsigned code = 0;
    int k = 0;
    for (int len = 1; len <= 16; len++) {
        for (int i = 0; i < counts[len - 1]; i++, k++) {
            if (k >= nsym || code >= (1u << len)) return JO_ERR_FORMAT;
            unsigned first = code << (16 - len), span = 1u << (16 - len);
            for (unsigned j = 0; j < span; j++) t->look[first + j] = (uint16_t)((len << 8) | symbols[k]);
            code++;
        }
        code <<= 1;
    }
    t->present = 1;
    return JO_OK;
}

static int rd16(const uint8_t *p) { return (p[0] << 8) | p[1]; }

static void free_decoder(Decoder *d) {
    for (int i = 0; i < 4; i++) { free(d->dc[i]); free(d->ac[i]); }
    for (int i = 0; i < 3; i++) { free(d->comp[i].coef); free(d->comp[i].plane); }
}

static int parse_headers(Decoder *d, const uint8_t *data, size_t len) {
    memset(d, 0, sizeof *d);
    if (len < 4 || data[0] != 0xFF || data[1] != 0xD8) return JO_ERR_FORMAT;
    size_t i = 2;
    int seen_sof = 0;
    int saw_jfif = 0, saw_adobe = 0, adobe_transform = 0;
    for (;;) {
        if (i + 4 > len) return JO_ERR_FORMAT;
        if (data[i] != 0xFF) return JO_ERR_FORMAT;
        while (i < len && data[i] == 0xFF) i++;          /* fill bytes */
        if (i >= len) return JO_ERR_FORMAT;
        int m = data[i++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (m == 0xD9) return JO_ERR_FORMAT;
        if (i + 2 > len) return JO_ERR_FORMAT;
        int L = rd16(data + i);
        if (L < 2 || i + (size_t)L > len) return JO_ERR_FORMAT;
        const uint8_t *seg = data + i + 2;
        int n = L - 2;
        if (m == 0xDB) {                                   /* DQT */
            int o = 0;
            while (o < n) {
                int pq = seg[o] >> 4, tq = seg[o] & 15;
                o++;
                if (tq > 3 || pq > 1 || o + 64 * (pq + 1) > n) return JO_ERR_FORMAT;
                for (int k = 0; k < 64; k++) {
                    d->quant[tq][ZIGZAG[k]] = (uint16_t)(pq ? rd16(seg + o + 2 * k) : seg[o + k]);
                }
                d->quant_present[tq] = 1;
                o += 64 * (pq + 1);
            }
        } else if (m == 0xC4) {                            /* DHT */
            int o = 0;
            while (o < n) {
                if (o + 17 > n) return JO_ERR_FORMAT;
                int tc = seg[o] >> 4, th = seg[o] & 15;
                if (tc > 1 || th > 3) return JO_ERR_FORMAT;
                int total = 0;
                for (int k = 0; k < 16; k++) total += seg[o + 1 + k];
                if (total > 256 || o + 17 + total > n) return JO_ERR_FORMAT;
                HuffTable **slot = tc ? &d->ac[th] : &d->dc[th];
                if (!*slot) *slot = (HuffTable *)malloc(sizeof(HuffTable));
                if (!*slot) return JO_ERR_SPACE;
                int rc = build_huff(*slot, seg + o + 1, seg + o + 17, total);
                if (rc) return rc;
                o += 17 + total;
            }
        } else if (m == 0xC0 || m == 0xC1) {               /* SOF0 / SOF1 */
            if (n < 6 || seen_sof) return JO_ERR_FORMAT;
            if (seg[0] != 8) return JO_ERR_UNSUPPORTED;
            d->h = rd16(seg + 1);
            d->w = rd16(seg + 3);
            d->ncomp = seg[5];
            if (d->w <= 0 || d->h <= 0) return JO_ERR_UNSUPPORTED;
            if (d->ncomp != 1 && d->ncomp != 3) return JO_ERR_UNSUPPORTED;
            if (n < 6 + 3 * d->ncomp) return JO_ERR_FORMAT;
            for (int c = 0; c < d->ncomp; c++) {
                d->comp[c].id = seg[6 + 3 * c];
                d->comp[c].hs = seg[7 + 3 * c] >> 4;
                d->comp[c].vs = seg[7 + 3 * c] & 15;
                d->comp[c].tq = seg[8 + 3 * c];
                if (d->comp[c].tq > 3) return JO_ERR_FORMAT;
            }
            seen_sof = 1;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC8 && m != 0xCC) {
            return JO_ERR_UNSUPPORTED;                      /* progressive, lossless, arithmetic */
        } else if (m == 0xDD) {                            /* DRI */
            if (n < 2) return JO_ERR_FORMAT;
            d->restart_interval = rd16(seg);
        } else if (m == 0xE0) {                            /* APP0: jdmarker.c examine_app0 */
            if (n >= 14 && !memcmp(seg, "JFIF\0", 5)) saw_jfif = 1;
        } else if (m == 0xEE) {                            /* APP14: jdmarker.c examine_app14 */
            if (n >= 12 && !memcmp(seg, "Adobe", 5)) { saw_adobe = 1; adobe_transform = seg[11]; }
        } else if (m == 0xDA) {                            /* SOS */
            if (!seen_sof) return JO_ERR_FORMAT;
            if (n < 1 || seg[0] != d->ncomp || n < 1 + 2 * d->ncomp + 3) return JO_ERR_UNSUPPORTED;
            for (int c = 0; c < d->ncomp; c++) {
                if (seg[1 + 2 * c] != d->comp[c].id) return JO_ERR_UNSUPPORTED;
                d->comp[c].td = seg[2 + 2 * c] >> 4;
                d->comp[c].ta = seg[2 + 2 * c] & 15;
                if (d->comp[c].td > 3 || d->comp[c].ta > 3) return JO_ERR_FORMAT;
            }
            d->scan = data + i + L;
            d->scan_len = len - (i + (size_t)L);
            break;
        }
        i += (size_t)L;
    }
    /* colour space of a 3-component file (libjpeg jdapimin.c default_decompress_parms): JFIF means
     * YCbCr; else an Adobe marker decides (transform 0 = RGB); else the component ids 'R','G','B'
     * mean RGB and anything else YCbCr */
    if (d->ncomp == 3) {
        if (saw_jfif) d->rgb = 0;
        else if (saw_adobe) d->rgb = adobe_transform == 0;
        else d->rgb = d->comp[0].id == 82 && d->comp[1].id == 71 && d->comp[2].id == 66;
    }
    /* geometry (libjpeg jdinput.c initial_setup / per_scan_setup) */
    if (d->ncomp == 1) {
        d->comp[0].hs = d->comp[0].vs = 1;                 /* a single-component scan is never interleaved */
    } else {
        int hs = d->comp[0].hs, vs = d->comp[0].vs;
        if (!((hs == 1 && vs == 1) || (hs == 2 && vs == 1) || (hs == 2 && vs == 2))) return JO_ERR_UNSUPPORTED;
        for (int c = 1; c < 3; c++)
            if (d->comp[c].hs != 1 || d->comp[c].vs != 1) return JO_ERR_UNSUPPORTED;
    }
    d->max_h = d->comp[0].hs;
    d->max_v = d->comp[0].vs;
    d->mcus_x = (d->w + 8 * d->max_h - 1) / (8 * d->max_h);
    d->mcus_y = (d->h + 8 * d->max_v - 1) / (8 * d->max_v);
    for (int c = 0; c < d->ncomp; c++) {
        Component *k = &d->comp[c];
        if (!d->quant_present[k->tq] || !d->dc[k->td] || !d->ac[k->ta]) return JO_ERR_FORMAT;
        k->blocks_w = d->mcus_x * k->hs;
        k->blocks_h = d->mcus_y * k->vs;
        k->ds_w = (d->w * k->hs + d->max_h - 1) / d->max_h;
        k->ds_h = (d->h * k->vs + d->max_v - 1) / d->max_v;
    }
    return JO_OK;
}

/* ---- entropy decoding (T.81 F.2; libjpeg jdhuff.c decode_mcu) ---- */
static int decode_scan(Decoder *d) {
    for (int c = 0; c < d->ncomp; c++) {
        Component *k = &d->comp[c];
        k->coef = (int16_t *)calloc((size_t)k->blocks_w * k->blocks_h * 64, sizeof(int16_t));
        if (!k->coef) return JO_ERR_SPACE;
    }
    BitReader br = {d->scan, d->scan + d->scan_len, 0, 0, 0, 0, 0};
    int pred[3] = {0, 0, 0};
    int todo = d->restart_interval;
    for (int my = 0; my < d->mcus_y; my++) {
        for (int mx = 0; mx < d->mcus_x; mx++) {
            if (d->restart_interval && todo == 0) {        /* process_restart */
                br.acc = 0; br.bits = 0; br.real = 0;      /* discard the partial byte */
                if (br.hit_marker || (br.p + 1 < br.end && br.p[0] == 0xFF)) {
                    while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) br.p++;
                    if (br.p + 1 < br.end) br.p += 2;
                    br.hit_marker = 0;
                    br.insufficient = 0;                   /* process_restart: a found marker re-arms decoding */
                }
                pred[0] = pred[1] = pred[2] = 0;
                todo = d->restart_interval;
            }
            /* decode_mcu: once the data has run dry the remaining MCUs are left as they are (all zero:
             * mid-grey after the IDCT); the MCU in which it happened is finished on zero bits */
            const int dry = br.insufficient;
            for (int c = 0; c < d->ncomp && !dry; c++) {
                Component *k = &d->comp[c];
                const HuffTable *dct = d->dc[k->td], *act = d->ac[k->ta];
                for (int by = 0; by < k->vs; by++) {
                    for (int bx = 0; bx < k->hs; bx++) {
                        int16_t *blk = k->coef + ((size_t)(my * k->vs + by) * k->blocks_w + (mx * k->hs + bx)) * 64;
                        int s = decode_symbol(&br, dct);
                        if (s) pred[c] += extend(br_get(&br, s), s);
                        blk[0] = (int16_t)pred[c];
                        for (int kk = 1; kk < 64; kk++) {
                            int rs = decode_symbol(&br, act);
                            int r = rs >> 4;
                            s = rs & 15;
                            if (s) {
                                kk += r;
                                int v = extend(br_get(&br, s), s);
                                if (kk < 64) blk[ZIGZAG[kk]] = (int16_t)v;
                            } else {
                                if (r != 15) break;
                                kk += 15;
                            }
                        }
                    }
                }
            }
            todo--;
        }
    }
    return JO_OK;
}

/* ---- dequantisation + inverse DCT (libjpeg jidctint.c, jpeg_idct_islow) ---- */
#define CONST_BITS 13
#define PASS1_BITS 2
#define FIX_0_298631336 2446
#define FIX_0_390180644 3196
#define FIX_0_541196100 4433
#define FIX_0_765366865 6270
#define FIX_0_899976223 7373
#define FIX_1_175875602 9633
#define FIX_1_501321110 12299
#define FIX_1_847759065 15137
#define FIX_1_961570560 16069
#define FIX_2_053119869 16819
#define FIX_2_562915447 20995
#define FIX_3_072711026 25172
#define DESCALE(x, n) (((x) + ((int32_t)1 << ((n) - 1))) >> (n))

/* One 8-point pass.  in[] with stride; writes 8 values descaled by `shift`. */
static void idct_1d(const int32_t *in, int stride, int32_t *out, int ostride, int shift) {
    int32_t z2 = in[2 * stride], z3 = in[6 * stride];
    int32_t z1 = (z2 + z3) * FIX_0_541196100;
    int32_t tmp2 = z1 + z3 * (-FIX_1_847759065);
    int32_t tmp3 = z1 + z2 * FIX_0_765366865;
    z2 = in[0];
    z3 = in[4 * stride];
    int32_t tmp0 = (int32_t)((uint32_t)(z2 + z3) << CONST_BITS);
    int32_t tmp1 = (int32_t)((uint32_t)(z2 - z3) << CONST_BITS);
    int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = in[7 * stride];
    tmp1 = in[5 * stride];
    tmp2 = in[3 * stride];
    tmp3 = in[1 * stride];
    z1 = tmp0 + tmp3;
    z2 = tmp1 + tmp2;
    z3 = tmp0 + tmp2;
    int32_t z4 = tmp1 + tmp3;
    int32_t z5 = (z3 + z4) * FIX_1_175875602;
    tmp0 *= FIX_0_298631336;
    tmp1 *= FIX_2_053119869;
    tmp2 *= FIX_3_072711026;
    tmp3 *= FIX_1_501321110;
    z1 *= -FIX_0_899976223;
    z2 *= -FIX_2_562915447;
    z3 *= -FIX_1_961570560;
    z4 *= -FIX_0_390180644;
    z3 += z5;
    z4 += z5;
    tmp0 += z1 + z3;
    tmp1 += z2 + z4;
    tmp2 += z2 + z3;
    tmp3 += z1 + z4;
    out[0 * ostride] = DESCALE(tmp10 + tmp3, shift);
    out[7 * ostride] = DESCALE(tmp10 - tmp3, shift);
    out[1 * ostride] = DESCALE(tmp11 + tmp2, shift);
    out[6 * ostride] = DESCALE(tmp11 - tmp2, shift);
    out[2 * ostride] = DESCALE(tmp12 + tmp1, shift);
    out[5 * ostride] = DESCALE(tmp12 - tmp1, shift);
    out[3 * ostride] = DESCALE(tmp13 + tmp0, shift);
    out[4 * ostride] = DESCALE(tmp13 - tmp0, shift);
}

static inline uint8_t clamp255(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

static void idct_block(const int16_t *coef, const uint16_t *q, uint8_t *dst, int dst_stride) {
    int32_t in[64], ws[64], row[8];
    for (int i = 0; i < 64; i++) in[i] = (int32_t)coef[i] * (int32_t)q[i];
    for (int c = 0; c < 8; c++) idct_1d(in + c, 8, ws + c, 8, CONST_BITS - PASS1_BITS);
    for (int r = 0; r < 8; r++) {
        idct_1d(ws + 8 * r, 1, row, 1, CONST_BITS + PASS1_BITS + 3);
        /* range_limit[(x) & RANGE_MASK] of jidctint.c == clamp(x + 128) for every x a conforming
         * stream can produce; the SIMD kernels cv2 runs saturate, as here. */
        for (int c = 0; c < 8; c++) dst[r * dst_stride + c] = clamp255(row[c] + 128);
    }
}

static int idct_all(Decoder *d) {
    for (int c = 0; c < d->ncomp; c++) {
        Component *k = &d->comp[c];
        int stride = k->blocks_w * 8;
        k->plane = (uint8_t *)malloc((size_t)stride * k->blocks_h * 8);
        if (!k->plane) return JO_ERR_SPACE;
        for (int by = 0; by < k->blocks_h; by++)
            for (int bx = 0; bx < k->blocks_w; bx++)
                idct_block(k->coef + ((size_t)by * k->blocks_w + bx) * 64, d->quant[k->tq],
                           k->plane + (size_t)by * 8 * stride + bx * 8, stride);
    }
    return JO_OK;
}

/* ---- chroma upsampling (libjpeg jdsample.c; context rows: jdmainct.c) ---- */
/* Value of the upsampled chroma plane at output pixel (x, y). */
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }

static int upsampled(const Decoder *d, const Component *k, int x, int y) {
    int stride = k->blocks_w * 8;
    const uint8_t *p = k->plane;
    if (d->max_h == 1) return p[(size_t)y * stride + x];                   /* fullsize_upsample */
    int cx = x >> 1, last = k->ds_w - 1;
    if (k->ds_w <= 2)                                                       /* jinit_upsampler: fancy only when */
        return p[(size_t)(d->max_v == 2 ? y >> 1 : y) * stride + cx];       /* downsampled_width > 2; else replicate */
    if (d->max_v == 1) {                                                    /* h2v1_fancy_upsample */
        const uint8_t *r = p + (size_t)y * stride;
        int v = r[cx];
        if (x & 1) return cx == last ? v : (3 * v + r[cx + 1] + 2) >> 2;
        return cx == 0 ? v : (3 * v + r[cx - 1] + 1) >> 2;
    }
    /* h2v2_fancy_upsample: nearer row weighs 3, farther row 1; then the same horizontally */
    int cy = y >> 1;
    int oy = clampi((y & 1) ? cy + 1 : cy - 1, 0, k->ds_h - 1);
    const uint8_t *r0 = p + (size_t)cy * stride, *r1 = p + (size_t)oy * stride;
    int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) {
        if (cx == last) return (4 * cur + 7) >> 4;
        return (3 * cur + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    }
    if (cx == 0) return (4 * cur + 8) >> 4;
    return (3 * cur + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}

/* ---- colour conversion (libjpeg jdcolor.c build_ycc_rgb_table / ycc_rgb_convert) ---- */
#define SCALEBITS 16
#define ONE_HALF ((int32_t)1 << (SCALEBITS - 1))
#define FIX(x) ((int32_t)((x) * (1L << SCALEBITS) + 0.5))

int jo_decode_bgr(const uint8_t *data, size_t len, uint8_t *out, size_t out_cap, int *w, int *h) {
    Decoder d;
    int rc = parse_headers(&d, data, len);
    if (rc) { free_decoder(&d); return rc; }
    *w = d.w;
    *h = d.h;
    if (!out) { free_decoder(&d); return JO_OK; }
    if (out_cap < (size_t)d.w * d.h * 3) { free_decoder(&d); return JO_ERR_SPACE; }
    rc = decode_scan(&d);
    if (!rc) rc = idct_all(&d);
    if (rc) { free_decoder(&d); return rc; }
    for (int y = 0; y < d.h; y++) {
        for (int x = 0; x < d.w; x++) {
            uint8_t *o = out + ((size_t)y * d.w + x) * 3;
            int yy = d.comp[0].plane[(size_t)y * d.comp[0].blocks_w * 8 + x];
            if (d.ncomp == 1) { o[0] = o[1] = o[2] = (uint8_t)yy; continue; }
            if (d.rgb) {                                   /* jdcolor.c rgb_rgb_convert: components are R, G, B */
                o[2] = (uint8_t)yy;
                o[1] = (uint8_t)upsampled(&d, &d.comp[1], x, y);
                o[0] = (uint8_t)upsampled(&d, &d.comp[2], x, y);
                continue;
            }
            int cb = upsampled(&d, &d.comp[1], x, y) - 128, cr = upsampled(&d, &d.comp[2], x, y) - 128;
            int r = yy + (int)((FIX(1.40200) * cr + ONE_HALF) >> SCALEBITS);
            int g = yy + (int)((-FIX(0.34414) * cb + ONE_HALF - FIX(0.71414) * cr) >> SCALEBITS);
            int b = yy + (int)((FIX(1.77200) * cb + ONE_HALF) >> SCALEBITS);
            o[0] = clamp255(b); o[1] = clamp255(g); o[2] = clamp255(r);
        }
    }
    free_decoder(&d);
    return JO_OK;
}

/* Quantised coefficients of component c in natural order, blocks in raster order over the
 * MCU-padded block grid (checks the product's entropy stage on its own). */
int jo_decode_coefficients(const uint8_t *data, size_t len, int c, int16_t *out, size_t out_cap,
                           int *blocks_w, int *blocks_h) {
    Decoder d;
    int rc = parse_headers(&d, data, len);
    if (!rc && (c < 0 || c >= d.ncomp)) rc = JO_ERR_FORMAT;
    if (rc) { free_decoder(&d); return rc; }
    *blocks_w = d.comp[c].blocks_w;
    *blocks_h = d.comp[c].blocks_h;
    size_t n = (size_t)d.comp[c].blocks_w * d.comp[c].blocks_h * 64;
    if (out) {
        if (out_cap < n) { free_decoder(&d); return JO_ERR_SPACE; }
        rc = decode_scan(&d);
        if (!rc) memcpy(out, d.comp[c].coef, n * sizeof(int16_t));
    }
    free_decoder(&d);
    return rc;
}
