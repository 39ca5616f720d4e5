// Host-side, one-time-per-params work: validation and the static per-dial tables.
//
// Everything here depends only on params.yml + the dial masks, never on a frame.  In the
// reference the corresponding values are recomputed per pixel per frame in Python
// (_reading.py:34-37 momentum terms, :55-66 per-pixel angle and squared distance,
// :89 sorted()); because they are functions of the pixel position and the dial centre
// only, they are tabulated once here with the SAME float64 operations (libm atan/pow,
// like CPython/numpy call), and the kernels look them up.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "melf_internal.h"

namespace {

// Python float %, for a positive modulus (Objects/floatobject.c: float_rem)
double py_mod_pos(double a, double m)
{
    double r = std::fmod(a, m);
    if (r != 0.0) {
        if (r < 0.0) r += m;
    } else {
        r = std::copysign(0.0, m);
    }
    return r;
}

// numpy float64 ** 2 is libm pow(x, 2.0); keep the call (the compiler would fold it)
double np_square(double x)
{
    volatile double two = 2.0;
    return std::pow(x, two);
}

}  // namespace

// get_angle_by_vector (_utils.py:18-42); NaN stands for None
double melf_host_angle_by_vector(double x, double y)
{
    if (y == 0.0) return x > 0.0 ? 0.25 : x < 0.0 ? 0.75 : NAN;
    double at = std::atan(x / y) / (2 * 3.141592653589793);
    return py_mod_pos(-at + (y > 0.0 ? 0.5 : 0.0), 1.0);
}

int melf_host_check_params(const melf_params *P, std::string *err)
{
    auto fail = [&](int code, const char *msg) { if (err) *err = msg; return code; };
    if (!P) return fail(MELF_ERR_INVALID, "params is NULL");
    if (P->struct_size != sizeof(melf_params))
        return fail(MELF_ERR_INVALID, "melf_params.struct_size mismatch (ABI)");
    // n_dials == 0: a locate-only context (melf_match_template with any template, _utils.py:91-97)
    if (P->n_dials < 0 || P->n_dials > MELF_MAX_DIALS)
        return fail(MELF_ERR_INVALID, "n_dials must be 0..MELF_MAX_DIALS");
    if (!P->tmpl || P->tmpl_w < 1 || P->tmpl_h < 1)
        return fail(MELF_ERR_INVALID, "missing dials template");
    if (P->tmpl_w > 256 || P->tmpl_h > 256)
        return fail(MELF_ERR_UNSUPPORTED, "dials template larger than 256x256");
    if (P->rect_x0 < 0 || P->rect_y0 < 0 || P->rect_x1 < P->rect_x0 || P->rect_y1 < P->rect_y0)
        return fail(MELF_ERR_UNSUPPORTED,
                    "meter_rect with negative or reversed coordinates (numpy would wrap)");
    if (P->hue_shift < 0 || P->hue_shift > 255)
        return fail(MELF_ERR_INVALID, "hue_shift must be 0..255");
    for (int d = 0; d < P->n_dials; d++) {
        const melf_dial &D = P->dials[d];
        if (!D.mask || !D.circle_mask) return fail(MELF_ERR_INVALID, "dial without masks");
        if (D.range_h < 0 || D.range_h > 255 || D.range_l < 0 || D.range_l > 255 ||
            D.range_s < 0 || D.range_s > 255)
            return fail(MELF_ERR_INVALID, "color_range channel outside 0..255");
        // get_dial_color reads dials_hls[y-2:y+3, x-2:x+3] with x,y = int(center)
        int px = (int)D.center_x - 2, py = (int)D.center_y - 2;
        if (!(D.center_x >= 0) || !(D.center_y >= 0) || px < 0 || py < 0 ||
            px + 5 > P->tmpl_w || py + 5 > P->tmpl_h)
            return fail(MELF_ERR_UNSUPPORTED,
                        "dial centre closer than 2 px to the template edge");
    }
    if (P->n_dials == 4)
        for (int k = 0; k < 4; k++)
            if (P->value_order[k] < 0 || P->value_order[k] >= 4)
                return fail(MELF_ERR_INVALID, "value_order entry outside 0..3");
    return MELF_OK;
}

int melf_host_build_dial(const melf_params *P, int d, MelfDialDev *out,
                         MelfRingTables *rings, std::string *err)
{
    const melf_dial &D = P->dials[d];
    const int tw = P->tmpl_w, th = P->tmpl_h;
    std::memset(out, 0, sizeof *out);

    int bx0 = tw, by0 = th, bx1 = -1, by1 = -1;
    for (int y = 0; y < th; y++)
        for (int x = 0; x < tw; x++)
            if (D.mask[(size_t)y * tw + x] | D.circle_mask[(size_t)y * tw + x]) {
                bx0 = std::min(bx0, x); bx1 = std::max(bx1, x);
                by0 = std::min(by0, y); by1 = std::max(by1, y);
            }
    if (bx1 < 0) { bx0 = by0 = 0; bx1 = by1 = 0; }   // empty mask: never any contour
    const int bw = bx1 - bx0 + 1, bh = by1 - by0 + 1;
    if (bw > MELF_DIAL_MAX_EXTENT || bh > MELF_DIAL_MAX_EXTENT) {
        if (err) *err = "dial mask larger than 60x60 pixels";
        return MELF_ERR_UNSUPPORTED;
    }
    out->win_x0 = bx0 - 2; out->win_y0 = by0 - 2;
    out->win_w = bw + 4;   out->win_h = bh + 4;
    out->patch_x = (int)D.center_x - 2;          // int() truncation, _reading.py:156
    out->patch_y = (int)D.center_y - 2;
    out->range_h = D.range_h; out->range_l = D.range_l; out->range_s = D.range_s;
    out->neg_mom = D.negative_momentum ? 1 : 0;
    out->aoz_turns = D.angle_of_zero / 360.0;

    for (int i = 0; i < 64; i++) {
        const int x = out->win_x0 + i, y = out->win_y0 + i;
        if (x >= 0 && x < tw) out->colvalid |= 1ull << i;
        if (y >= 0 && y < th) out->rowvalid |= 1ull << i;
        const double dx = (double)x - D.center_x, dy = (double)y - D.center_y;
        out->termx[i] = (dx < 0 ? -1 : 1) * np_square(dx);
        out->termy[i] = (dy < 0 ? -1 : 1) * np_square(dy);
    }
    struct Entry { double a, s; uint16_t pos; };
    std::vector<Entry> entries;
    for (int r = 0; r < 64; r++) {
        const int y = out->win_y0 + r;
        if (y < 0 || y >= th) continue;
        for (int c = 0; c < 64; c++) {
            const int x = out->win_x0 + c;
            if (x < 0 || x >= tw) continue;
            const size_t i = (size_t)y * tw + x;
            if (D.mask[i]) out->disc[r] |= 1ull << c;
            if (D.circle_mask[i]) {
                out->ring[r] |= 1ull << c;
                const double dx = (double)x - D.center_x, dy = (double)y - D.center_y;
                const double a = melf_host_angle_by_vector(dx, dy);
                if (a != a) continue;            // None: never kept (_reading.py:61)
                entries.push_back({a, np_square(dx) + np_square(dy), (uint16_t)((r << 8) | c)});
            }
        }
    }
    // needed inRange pixels: 2-px Chebyshev neighbourhood of disc | ring inside the rect
    {
        uint64_t dr[64], need[64];
        for (int r = 0; r < 64; r++) dr[r] = out->disc[r] | out->ring[r];
        for (int r = 0; r < 64; r++) {
            uint64_t v = 0;
            for (int k = -2; k <= 2; k++)
                if (r + k >= 0 && r + k < 64) v |= dr[r + k];
            v = v | (v << 1) | (v << 2) | (v >> 1) | (v >> 2);
            need[r] = ((out->rowvalid >> r) & 1) ? (v & out->colvalid) : 0;
        }
        out->need_off = (int32_t)rings->need.size();
        int base = 0;
        for (int r = 0; r < 64; r++) {
            out->row_base[r] = (uint16_t)base;
            out->row_c0[r] = 0;
            if (!need[r]) continue;
            int c0 = 0, c1 = 63;
            while (!((need[r] >> c0) & 1)) c0++;
            while (!((need[r] >> c1) & 1)) c1--;
            out->row_c0[r] = (uint8_t)c0;
            for (int c = c0; c <= c1; c++) rings->need.push_back((uint16_t)((r << 8) | c));
            base += c1 - c0 + 1;
        }
        out->row_base[64] = (uint16_t)base;
        out->n_need = base;
    }
    std::stable_sort(entries.begin(), entries.end(), [](const Entry &p, const Entry &q) {
        if (p.a != q.a) return p.a < q.a;
        return p.s < q.s;
    });
    out->ring_off = (int32_t)rings->pos.size();
    out->n_ring = (int32_t)entries.size();
    for (const Entry &e : entries) {
        rings->pos.push_back(e.pos);
        rings->angle.push_back(e.a);
        rings->sqd.push_back(e.s);
    }
    return MELF_OK;
}

// ---- host-only debug export (tests/test_host_tables.py; needs no GPU) ----------------
extern "C" int melf_host_dial_tables(const melf_params *P, int d,
                                     int32_t *geom /*[6]: win_x0,win_y0,win_w,win_h,patch_x,patch_y*/,
                                     uint64_t *words /*[2 + 2*64]: colvalid,rowvalid,disc,ring*/,
                                     double *terms /*[128]: termx, termy*/,
                                     int32_t *n_ring, uint16_t *ring_pos, double *ring_angle,
                                     double *ring_sqd, int32_t ring_cap)
{
    std::string err;
    int rc = melf_host_check_params(P, &err);
    if (rc) return rc;
    if (d < 0 || d >= P->n_dials) return MELF_ERR_INVALID;
    MelfDialDev dev;
    MelfRingTables rings;
    rc = melf_host_build_dial(P, d, &dev, &rings, &err);
    if (rc) return rc;
    geom[0] = dev.win_x0; geom[1] = dev.win_y0; geom[2] = dev.win_w; geom[3] = dev.win_h;
    geom[4] = dev.patch_x; geom[5] = dev.patch_y;
    words[0] = dev.colvalid; words[1] = dev.rowvalid;
    std::memcpy(words + 2, dev.disc, sizeof dev.disc);
    std::memcpy(words + 2 + 64, dev.ring, sizeof dev.ring);
    std::memcpy(terms, dev.termx, sizeof dev.termx);
    std::memcpy(terms + 64, dev.termy, sizeof dev.termy);
    *n_ring = dev.n_ring;
    if (dev.n_ring > ring_cap) return MELF_ERR_INVALID;
    std::memcpy(ring_pos, rings.pos.data(), dev.n_ring * sizeof(uint16_t));
    std::memcpy(ring_angle, rings.angle.data(), dev.n_ring * sizeof(double));
    std::memcpy(ring_sqd, rings.sqd.data(), dev.n_ring * sizeof(double));
    return MELF_OK;
}

// ---------------------------------------------------------------------------------------------
// Host-side batch helpers of the ingest row: what get_meter_values does per file
// before any pixel exists -- open + read (cv2.imread's file access, _image.py:49) and the header
// probe that decides whether the GPU decoder may take the stream -- for a whole batch of files on a
// few host threads, so that the Python layer spends microseconds per batch, not per file.
// ---------------------------------------------------------------------------------------------
#include <cstdio>
#include <thread>
#include <sys/stat.h>

#include "melf_jpeg.h"

namespace {

template <class F>
void parallel_for(int n, int max_threads, F body)
{
    int nt = (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > max_threads) nt = max_threads;
    if (nt > (n + 63) / 64) nt = (n + 63) / 64;
    if (nt <= 1) { for (int i = 0; i < n; i++) body(i); return; }
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; t++)
        pool.emplace_back([=] { for (int i = t; i < n; i += nt) body(i); });
    for (std::thread &th : pool) th.join();
}

// EXIF orientation tag (0x0112) of a JPEG stream, 1 when absent: cv2.imread rotates by it, the ingest
// kernels do not, so anything but 1 stays on the host (same walk as _image.exif_orientation)
int exif_orientation(const uint8_t *d, size_t len)
{
    size_t pos = 2;
    while (pos + 4 <= len && d[pos] == 0xFF) {
        const int marker = d[pos + 1];
        if (marker == 0xDA || marker == 0xD9) break;
        const size_t size = ((size_t)d[pos + 2] << 8) | d[pos + 3];
        if (size < 2) break;
        if (marker == 0xE1 && pos + 2 + size <= len && size >= 2 + 14 && !std::memcmp(d + pos + 4, "Exif\0\0", 6)) {
            const uint8_t *t = d + pos + 10;
            const size_t tl = size - 8;
            const bool le = t[0] == 'I' && t[1] == 'I';
            auto rd = [&](size_t o, int nb) -> uint32_t {
                uint32_t v = 0;
                for (int k = 0; k < nb; k++) v |= (uint32_t)t[o + (le ? k : nb - 1 - k)] << (8 * k);
                return v;
            };
            const size_t ifd = rd(4, 4);
            if (ifd + 2 > tl) return 1;
            const int entries = (int)rd(ifd, 2);
            for (int k = 0; k < entries; k++) {
                const size_t e = ifd + 2 + 12 * (size_t)k;
                if (e + 12 > tl) break;
                if (rd(e, 2) == 0x0112) return (int)rd(e + 8, 2);
            }
            return 1;
        }
        pos += 2 + size;
    }
    return 1;
}

}  // namespace

extern "C" {

// sizes[i] = size of paths[i] in bytes, or -1 when it cannot be opened (cv2.imread would return None)
int melf_file_sizes(const char *const *paths, int n, int64_t *sizes)
{
    if (n < 0 || (n > 0 && (!paths || !sizes))) return MELF_ERR_INVALID;
    parallel_for(n, 8, [&](int i) {
        struct stat st;
        sizes[i] = (stat(paths[i], &st) == 0 && S_ISREG(st.st_mode)) ? (int64_t)st.st_size : -1;
    });
    return MELF_OK;
}

// file i -> blob[offsets[i] .. offsets[i+1]); ok[i] = 0 when the file could not be read in full
int melf_read_files(const char *const *paths, int n, uint8_t *blob, const uint64_t *offsets, int32_t *ok)
{
    if (n < 0 || (n > 0 && (!paths || !blob || !offsets || !ok))) return MELF_ERR_INVALID;
    parallel_for(n, 16, [&](int i) {
        const size_t want = (size_t)(offsets[i + 1] - offsets[i]);
        ok[i] = 0;
        if (FILE *fp = std::fopen(paths[i], "rb")) {
            ok[i] = std::fread(blob + offsets[i], 1, want, fp) == want ? 1 : 0;
            std::fclose(fp);
        }
    });
    return MELF_OK;
}

// Header probe of n streams: info[8 i ..] = {h, w, components, h_samp, v_samp, EXIF orientation,
// end-of-image marker within the last 64 bytes (0/1), code}; code 0 = the GPU decoder takes the stream,
// else the MELF_ERR_* melf_jpeg_info would return (the file then takes the cv2.imread route).
int melf_jpeg_probe_batch(const uint8_t *jpeg, const uint64_t *offsets, int n, int32_t *info)
{
    if (n < 0 || (n > 0 && (!jpeg || !offsets || !info))) return MELF_ERR_INVALID;
    parallel_for(n, 8, [&](int i) {
        const uint8_t *d = jpeg + offsets[i];
        const size_t len = (size_t)(offsets[i + 1] - offsets[i]);
        int32_t *o = info + 8 * (size_t)i;
        int h = 0, w = 0, nc = 0, hs = 0, vs = 0;
        o[7] = (len >= 2 && d[0] == 0xFF && d[1] == 0xD8) ? melf_jpeg_probe(d, len, &h, &w, &nc, &hs, &vs, nullptr)
                                                          : MELF_ERR_INVALID;
        o[0] = h; o[1] = w; o[2] = nc; o[3] = hs; o[4] = vs;
        o[5] = o[7] == 0 ? exif_orientation(d, len) : 1;
        int eoi = 0;
        for (size_t k = len >= 64 ? len - 64 : 0; k + 1 < len; k++)
            if (d[k] == 0xFF && d[k + 1] == 0xD9) { eoi = 1; break; }
        o[6] = eoi;
    });
    return MELF_OK;
}

}  // extern "C"
