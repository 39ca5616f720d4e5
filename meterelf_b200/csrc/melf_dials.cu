// Per-dial needle reading: one warp per (frame, dial), one block per frame.
//
// Replaces, for a whole batch, the per-dial loop of get_meter_value (_reading.py:28-96),
// get_needle_points (_reading.py:118-151), get_dial_color (_reading.py:154-160),
// get_mask_by_color (_utils.py:113-119) and determine_value_by_dial_positions
// (_reading.py:163-182).
//
// Formulation (DESIGN.md "dial kernel"): everything the reference does on the 119x188
// dials rect is ANDed with the dial's disc mask, so only the disc's bounding box plus a
// 2-pixel halo matters.  That window (<= 64x64) is held as one 64-bit word per row, two
// rows per lane, entirely in registers:
//   inRange       -> ballots of the per-pixel HLS test (HLS recomputed from BGR on the fly)
//   dilate/erode  -> shifts and ANDs/ORs of row words, neighbours by warp shuffle
//   findContours + contourArea + drawContours(FILLED)
//                 -> bit-parallel 8-connected flood per component, 4-connected exterior
//                    flood for hole filling, bit-quad area (2*#quads4 + #quads3)
//   momentum, ring angles, trimmed weighted mean
//                 -> sequential float64 sums in the reference's order over host-built
//                    static tables (per-pixel angle / squared distance / sort order depend
//                    only on params), so dial positions are bit-identical to the oracle.
#include "melf_common.cuh"
#ifndef MELF_DIALS_U4
#define MELF_DIALS_U4 2      // unroll factors of the kernel's loops (code size against loop overhead: the kernel is
#define MELF_DIALS_U2 1      // ~75 KB of SASS and 18 % of its stall samples are instruction fetch)
#endif
#define MELF_PRAGMA_(x) _Pragma(#x)
#define MELF_UNROLL(n) MELF_PRAGMA_(unroll n)
#include "melf_args.h"

namespace {

constexpr u64 ALL1 = ~0ull;

__device__ __forceinline__ u64 shfl_u64(u64 v, int src)
{
    return __shfl_sync(0xffffffffu, v, src);
}
__device__ __forceinline__ u64 shfl_up1(u64 v, u64 fill, int lane)
{
    u64 r = __shfl_up_sync(0xffffffffu, v, 1);
    return lane == 0 ? fill : r;
}
__device__ __forceinline__ u64 shfl_down1(u64 v, u64 fill, int lane)
{
    u64 r = __shfl_down_sync(0xffffffffu, v, 1);
    return lane == 31 ? fill : r;
}
__device__ __forceinline__ u64 warp_or64(u64 v)
{
    uint32_t lo = __reduce_or_sync(0xffffffffu, (uint32_t)v);
    uint32_t hi = __reduce_or_sync(0xffffffffu, (uint32_t)(v >> 32));
    return ((u64)hi << 32) | lo;
}

// fill every run of ones of m that contains a bit of s (carry propagation both ways)
__device__ __forceinline__ u64 spread(u64 s, u64 m)
{
    s &= m;
    const u64 up = (((m + s) ^ m) | s) & m;
    const u64 mr = __brevll(m), sr = __brevll(s);
    const u64 dn = __brevll((((mr + sr) ^ mr) | sr) & mr);
    return up | dn;
}
__device__ __forceinline__ u64 h3(u64 x) { return x | (x << 1) | (x >> 1); }
__device__ __forceinline__ u64 e3(u64 x) { return x & ((x << 1) | 1ull) & ((x >> 1) | (1ull << 63)); }

// 8-connected flood of the seed inside `allowed` (rows 2*lane = a, 2*lane+1 = b)
__device__ void flood8(u64 &ca, u64 &cb, const u64 al_a, const u64 al_b, const int lane)
{
    for (int it = 0; it < 8192; ++it) {
        u64 na = spread(ca, al_a), nb = spread(cb, al_b);
        const u64 up = shfl_up1(nb, 0, lane);      // row 2*lane-1
        const u64 dn = shfl_down1(na, 0, lane);    // row 2*lane+2
        na |= h3(up | nb) & al_a;
        nb |= h3(na | dn) & al_b;
        const bool changed = (na != ca) | (nb != cb);
        ca = na; cb = nb;
        if (!__any_sync(0xffffffffu, changed)) break;
    }
}

// component -> component plus its holes (background not 4-connected to the outside)
__device__ void fill_holes(const u64 ca, const u64 cb, u64 &fa, u64 &fb, const int lane)
{
    const uint32_t rows = __ballot_sync(0xffffffffu, (ca | cb) != 0);
    const int lmin = __ffs(rows) - 1, lmax = 31 - __clz(rows);
    const u64 uni = warp_or64(ca | cb);
    const int cmin = __ffsll((long long)uni) - 1, cmax = 63 - __clzll((long long)uni);
    const u64 colrange = (cmax == 63 ? ALL1 : ((1ull << (cmax + 1)) - 1)) & ~((1ull << cmin) - 1);
    const bool inrows = lane >= lmin && lane <= lmax;
    u64 ra = inrows ? ~colrange : ALL1, rb = ra;          // everything outside the bbox
    const u64 al_a = ~ca, al_b = ~cb;
    for (int it = 0; it < 8192; ++it) {
        u64 na = ra | spread(ra & al_a, al_a), nb = rb | spread(rb & al_b, al_b);
        const u64 up = shfl_up1(nb, ALL1, lane);
        const u64 dn = shfl_down1(na, ALL1, lane);
        na |= (up | nb) & al_a;
        nb |= (na | dn) & al_b;
        const bool changed = (na != ra) | (nb != rb);
        ra = na; rb = nb;
        if (!__any_sync(0xffffffffu, changed)) break;
    }
    fa = ~ra; fb = ~rb;
}

__device__ __forceinline__ int quad_area2(u64 x, u64 y)
{
    const u64 q1 = x >> 1, q3 = y >> 1;
    const u64 t01 = x & q1, x01 = x ^ q1, t23 = y & q3, x23 = y ^ q3;
    return 2 * __popcll(t01 & t23) + __popcll((t01 & x23) | (x01 & t23));
}

// twice the area of the polygon through the boundary pixel centres of the filled set
__device__ int filled_area2(const u64 fa, const u64 fb, const int lane)
{
    const u64 up = shfl_up1(fb, 0, lane);
    int a = quad_area2(up, fa) + quad_area2(fa, fb);
    if (lane == 31) a += quad_area2(fb, 0);
    return __reduce_add_sync(0xffffffffu, a);
}

// Python float % for a positive modulus.  Not inlined: fmod is ~110 instructions, and the kernel's code size costs
// time (18 % of its stall samples were instruction fetch with five inlined copies).
__device__ __noinline__ double py_mod_pos(double a, double m)
{
    double r = fmod(a, m);
    if (r != 0.0) {
        if (r < 0.0) r = __dadd_rn(r, m);
    } else {
        r = 0.0;
    }
    return r;
}

// Python float % 1.0: fmod(a, 1) is a - trunc(a), an exact subtraction (the zero result's sign does not matter, it
// becomes +0.0 as in py_mod_pos)
__device__ __forceinline__ double py_mod1(double a)
{
    double r = __dsub_rn(a, trunc(a));
    if (r != 0.0) {
        if (r < 0.0) r = __dadd_rn(r, 1.0);
    } else {
        r = 0.0;
    }
    return r;
}

// get_angle_by_vector (_utils.py:18-42); NaN = None
__device__ double angle_by_vector(double x, double y)
{
    if (y == 0.0) return x > 0.0 ? 0.25 : x < 0.0 ? 0.75 : __longlong_as_double(0x7ff8000000000000ll);
    const double at = __ddiv_rn(atan(__ddiv_rn(x, y)), 2 * 3.141592653589793);
    return py_mod1(__dadd_rn(-at, y > 0.0 ? 0.5 : 0.0));
}

__device__ __forceinline__ bool near_momentum(double a, double mom)
{
    const double diff = fabs(__dsub_rn(a, mom));
    const double alt = fabs(__dsub_rn(diff, 1.0));
    const double dist = alt < diff ? alt : diff;        // Python min()
    return dist < 0.25;                                 // false for NaN
}

__device__ __forceinline__ int carry_digit(double r, double lower)
{
    const double frac = py_mod1(r);
    int v = (int)r + ((frac > 0.55 && lower <= 2.0) ? 1 : 0) - ((frac < 0.45 && lower >= 8.0) ? 1 : 0);
    v %= 10;
    return v < 0 ? v + 10 : v;
}

struct DialOut {
    double pos;
    int flag;       // 0 read, 1 no contours, 2 unreadable
};

}  // namespace


extern __shared__ __align__(16) uint8_t melf_dials_dyn[];

// Work distribution: a CTA is DIALS_WPB warps that all read the SAME dial d of consecutive frames
// (blockIdx = frame group * n_dials + d), so the warps of a CTA do equal work -- with one CTA per
// frame the warps of the small dials idled 18 % of all warp samples at the closing barrier
// (profiles/r01_v7_ncu_dials.txt).  The frame result is assembled by whichever dial warp of
// the frame finishes last (MelfDialSlot).
constexpr int DIALS_WPB = 4;

__global__ void __launch_bounds__(32 * DIALS_WPB)
melf_dials_kernel(const MelfDialsArgs A)
{
    __shared__ u64 nm_s[DIALS_WPB][64];

    const int wslot = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int d = (int)(blockIdx.x % (unsigned)A.n_dials);
    const int f = (int)(blockIdx.x / (unsigned)A.n_dials) * DIALS_WPB + wslot;
    if (f >= A.n_frames) return;                         // no block-wide barrier below

    const u64 key = A.match_key[f];
    const float max_val = melf_key_score(key);
    const uint32_t midx = melf_key_index(key);
    const int mx = (int)(midx % (uint32_t)A.geom.res_w), my = (int)(midx / (uint32_t)A.geom.res_w);

    if ((double)max_val < A.threshold) {                 // _image.py:62-64 (uniform over the frame's warps)
        if (d == 0 && lane == 0) {
            melf_result r;
            r.status = MELF_ST_DIALS_NOT_FOUND;
            r.match_x = mx; r.match_y = my; r.max_val = max_val;
            r.failed_dial = -1; r.unreadable_mask = 0;
            for (int i = 0; i < MELF_MAX_DIALS; i++) r.dial_pos[i] = __longlong_as_double(0x7ff8000000000000ll);
            r.value = __longlong_as_double(0x7ff8000000000000ll);
            A.out[f] = r;
        }
        return;
    }

    long long *tl = (A.timeline && f < 8 && d == 0 && lane == 0) ? A.timeline + 8 * f : nullptr;
    if (tl) tl[0] = clock64();
    const MelfDialDev &D = A.dials[d];
    // pixel (u, v) of the dials rect -> crop column/row -> frame address
    const uint8_t *frame = A.bgr + (size_t)f * A.frame_stride;
    const int crop_w = A.geom.crop_w;
    auto hls_at = [&](int u, int v) -> uint32_t {
        const int cx = mx + u, cy = my + v;
        const uint8_t *p = frame + (size_t)(A.geom.crop_y0 + cy) * A.row_stride + 3 * (size_t)(A.geom.crop_x0 + cx);
        const int B = p[0], G = p[1], R = p[2];
        if (A.s_tail && melf_is_tail_col(cx, crop_w))       // OpenCV's scalar-tail S association
            return melf_hls_pixel(B, G, R, true, A.hue_shift);
        return melf_hls_arith((uint32_t)B, (uint32_t)G, (uint32_t)R, A.hue_shift);
    };

    const uint16_t *need = A.need_pos + D.need_off;      // flat list of the window pixels inRange visits
    const int n_need = D.n_need;
    uint32_t pos_nx = lane < n_need ? need[lane] : 0u;

    // ---- stage the window's BGR bytes in shared memory ----------------------------------
    // Row r of the window occupies 3 * win_w bytes starting at an arbitrary address: fetch the
    // 16-byte chunks that overlap its valid columns with cp.async (global -> shared without
    // registers, so every chunk of every row is in flight at once), keeping the chunk grid
    // aligned to global memory so that the byte of window column c sits at
    // pix[r * pitch + (rowaddr(r) & 15) + 3 * c].  The dial colour below is computed while
    // the copies fly.
    uint8_t *pix = melf_dials_dyn + (size_t)wslot * A.pix_bytes;
    const u64 V = D.colvalid, RV = D.rowvalid;
    const uint8_t *win0 = frame + (size_t)(A.geom.crop_y0 + my + D.win_y0) * A.row_stride +
                          3 * (ptrdiff_t)(A.geom.crop_x0 + mx + D.win_x0);
    {
        const int cmin = __ffsll((long long)V) - 1, cmax = 63 - __clzll((long long)V);
        const int n_chunk = A.pix_pitch >> 4;
        const int sub = lane >> 4, ch = lane & 15;
        const uint32_t pix_s = (uint32_t)__cvta_generic_to_shared(pix);
        const uintptr_t end = reinterpret_cast<uintptr_t>(A.bgr_end);
        if (ch < n_chunk && V) {
MELF_UNROLL(MELF_DIALS_U4)
            for (int rr = sub; rr < D.win_h; rr += 2) {
                if (!((RV >> rr) & 1)) continue;
                const uintptr_t rowp = reinterpret_cast<uintptr_t>(win0 + (size_t)rr * A.row_stride);
                const uintptr_t c_lo = (rowp & ~(uintptr_t)15) + 16 * (uintptr_t)ch;
                if (c_lo + 16 > rowp + 3 * (uintptr_t)cmin && c_lo < rowp + 3 * (uintptr_t)(cmax + 1)) {
                    if (c_lo + 16 <= end) {
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                                     :: "r"(pix_s + (uint32_t)(rr * A.pix_pitch + 16 * ch)), "l"(c_lo) : "memory");
                    } else {
                        // the chunk straddles the end of the frame memory: copy only what exists
                        for (uintptr_t q = c_lo; q < end; ++q)
                            pix[rr * A.pix_pitch + 16 * ch + (int)(q - c_lo)] = *reinterpret_cast<const uint8_t *>(q);
                    }
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

    // ---- get_dial_color: rounded mean of the 5x5 patch at int(centre) -----------------
    int sh = 0, sl = 0, ss = 0;
    if (lane < 25) {
        const uint32_t c = hls_at(D.patch_x + lane % 5, D.patch_y + lane / 5);
        sh = c & 255; sl = (c >> 8) & 255; ss = (c >> 16) & 255;
    }
    sh = __reduce_add_sync(0xffffffffu, sh);
    sl = __reduce_add_sync(0xffffffffu, sl);
    ss = __reduce_add_sync(0xffffffffu, ss);
    // round(sum / 25.0): 25 is odd so there are no ties
    const int col_h = (2 * sh + 25) / 50, col_l = (2 * sl + 25) / 50, col_s = (2 * ss + 25) / 50;
    const int lo_h = max(col_h - D.range_h, 0), hi_h = min(col_h + D.range_h, 255);
    const int lo_l = max(col_l - D.range_l, 0), hi_l = min(col_l + D.range_l, 255);
    const int lo_s = max(col_s - D.range_s, 0), hi_s = min(col_s + D.range_s, 255);

    if (tl) tl[1] = clock64();
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    auto hls_win = [&](int cc, int rr) -> uint32_t {
        const uint8_t *rowp = win0 + (size_t)rr * A.row_stride;
        const uint8_t *p = pix + rr * A.pix_pitch + (reinterpret_cast<uintptr_t>(rowp) & 15) + 3 * cc;
        const uint32_t B = p[0], G = p[1], R = p[2];
        const int cx = mx + D.win_x0 + cc;
        if (A.s_tail && melf_is_tail_col(cx, crop_w))       // OpenCV's scalar-tail S association
            return melf_hls_pixel((int)B, (int)G, (int)R, true, A.hue_shift);
        return melf_hls_arith(B, G, R, A.hue_shift);
    };

    if (tl) tl[7] = clock64() - tl[0];                   // debug: window staged
    // ---- inRange over the needed window pixels -> row words -----------------------------
    // Only pixels within 2 px of disc | ring can reach a bit of the closed mask that is used
    // (MelfDialDev::row_base): they are visited in a flat raster list, 32 per ballot, and the
    // ballot words are cut back into 64-bit row words afterwards.
    u64 m0a = 0, m0b = 0;
    {
        uint32_t *bal_s = reinterpret_cast<uint32_t *>(nm_s[wslot]);       // 128 words >= 4096 / 32
        const float s_lo_f = (float)(lo_s - 1), s_hi_f = (float)(hi_s + 1);
MELF_UNROLL(MELF_DIALS_U2)
        for (int base = 0; base < n_need; base += 32) {
            const int p = base + lane;
            const bool valid = p < n_need;
            // the list entry was fetched one iteration ahead (its L2 latency is off the critical path)
            const uint32_t pos = pos_nx;
            pos_nx = p + 32 < n_need ? need[p + 32] : 0u;
            // cheap tests first: when no pixel of the 32 can pass the lightness and saturation ranges
            // (dial face against a red needle: most groups), the saturation division and the hue are skipped
            const int cc = (int)(pos & 255u), rr = (int)(pos >> 8);
            const uint8_t *rowp = win0 + (size_t)rr * A.row_stride;
            const uint8_t *pp = pix + rr * A.pix_pitch + (reinterpret_cast<uintptr_t>(rowp) & 15) + 3 * cc;
            const uint32_t Bv = pp[0], Gv = pp[1], Rv = pp[2];
            const bool tail = A.s_tail && melf_is_tail_col(mx + D.win_x0 + cc, crop_w);
            const MelfHlsStage1 st1 = melf_hls_stage1(Bv, Gv, Rv);
            // conservative saturation pre-test without the division: S = rint(255 diff / den) can only lie
            // in [lo_s, hi_s] if 255 diff lies in [(lo_s - 1) den, (hi_s + 1) den] (half a unit of slack on
            // each side against float32 rounding errors of ~1e-7 relative)
            const float d255 = __fmul_rn(__fsub_rn(st1.vmax, st1.vmin), 255.f);
            const float den0 = fminf(st1.sm, __fsub_rn(2.f, st1.sm));
            const bool s_maybe = d255 >= __fmul_rn(s_lo_f, den0) && d255 <= __fmul_rn(s_hi_f, den0);
            const bool l_ok = valid && (int)st1.L >= lo_l && (int)st1.L <= hi_l && (s_maybe || tail);
            bool bit = false;
            if (__any_sync(0xffffffffu, l_ok)) {
                // OpenCV's scalar-tail columns differ in S only; L is the same in both forms
                const uint32_t c = tail ? melf_hls_pixel((int)Bv, (int)Gv, (int)Rv, true, A.hue_shift)
                                        : melf_hls_stage2(st1, A.hue_shift);
                const int H = c & 255, S = (c >> 16) & 255;
                bit = l_ok && H >= lo_h && H <= hi_h && S >= lo_s && S <= hi_s;
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            if (lane == 0) bal_s[base >> 5] = bal;
        }
        __syncwarp();
        auto row_word = [&](int r) -> u64 {
            const int b0 = D.row_base[r], n = D.row_base[r + 1] - b0;
            if (n <= 0) return 0ull;
            const int wi = b0 >> 5, sh = b0 & 31;
            const int last = (n_need - 1) >> 5;
            const u64 w0 = bal_s[wi], w1 = wi + 1 <= last ? bal_s[wi + 1] : 0u, w2 = wi + 2 <= last ? bal_s[wi + 2] : 0u;
            u64 v = ((w0 | (w1 << 32)) >> sh) | (sh ? (w2 << (64 - sh)) : 0ull);
            if (n < 64) v &= (1ull << n) - 1ull;
            return v << D.row_c0[r];
        };
        m0a = row_word(2 * lane);
        m0b = row_word(2 * lane + 1);
        __syncwarp();                                       // bal_s (= nm_s[wslot]) is reused by the ring phase
    }

    if (tl) tl[2] = clock64();
    // ---- 3x3 dilate then erode (pixels outside the dials rect do not take part) -------
    const bool rva = (RV >> (2 * lane)) & 1, rvb = (RV >> (2 * lane + 1)) & 1;
    u64 era, erb;
    {
        const u64 ha = h3(m0a), hb = h3(m0b);
        const u64 hup = shfl_up1(hb, 0, lane), hdn = shfl_down1(ha, 0, lane);
        const u64 da = rva ? ((hup | ha | hb) & V) : 0;
        const u64 db = rvb ? ((ha | hb | hdn) & V) : 0;
        const u64 pa = rva ? e3(da | ~V) : ALL1;
        const u64 pb = rvb ? e3(db | ~V) : ALL1;
        const u64 pup = shfl_up1(pb, ALL1, lane), pdn = shfl_down1(pa, ALL1, lane);
        era = rva ? ((pup & pa & pb) & V) : 0;
        erb = rvb ? ((pa & pb & pdn) & V) : 0;
    }

    if (tl) tl[3] = clock64();
    // ---- components: largest filled area wins, first discovered wins ties -------------
    u64 rem_a = era & D.disc[2 * lane], rem_b = erb & D.disc[2 * lane + 1];
    u64 best_a = 0, best_b = 0;
    int best_area2 = -1, ncomp = 0;
    for (int guard = 0; guard < 4096; ++guard) {
        const uint32_t rows = __ballot_sync(0xffffffffu, (rem_a | rem_b) != 0);
        if (!rows) break;
        const int l0 = __ffs(rows) - 1;
        u64 ca = 0, cb = 0;
        if (lane == l0) {
            if (rem_a) ca = rem_a & (0 - rem_a); else cb = rem_b & (0 - rem_b);
        }
        flood8(ca, cb, rem_a, rem_b, lane);
        u64 fa, fb;
        fill_holes(ca, cb, fa, fb, lane);
        const int a2 = filled_area2(fa, fb, lane);
        ++ncomp;
        if (a2 > best_area2) { best_area2 = a2; best_a = fa; best_b = fb; }
        rem_a &= ~fa; rem_b &= ~fb;
    }

    if (tl) tl[4] = clock64();
    DialOut res;
    res.pos = __longlong_as_double(0x7ff8000000000000ll);
    res.flag = 0;
    int n_needle = 0, n_kept = 0;
    double mom_x = 0.0, mom_y = 0.0, mom_angle = res.pos;
    u64 nm_a = 0, nm_b = 0;

    if (best_area2 < 0) {
        res.flag = 1;                                    // _reading.py:137-138
    } else {
        // _reading.py:140-148: filled contour if area > 100 else the closed mask itself
        nm_a = best_area2 > 200 ? best_a : era;
        nm_b = best_area2 > 200 ? best_b : erb;

        // ---- momentum (_reading.py:32-41): sequential sums in raster order -----------
        // Per row: the lanes compact the x terms of the set pixels into a small shared buffer
        // (lane j owns columns j and j + 32), then every lane runs the same dependent chain of
        // float64 additions over it, in the reference's raster order.  The loads of the chain
        // do not depend on the sums, so only the DADD latency remains on the critical path.
        const u64 np_a = nm_a & D.disc[2 * lane], np_b = nm_b & D.disc[2 * lane + 1];
        n_needle = __reduce_add_sync(0xffffffffu, __popcll(np_a) + __popcll(np_b));
        {
            double *tbuf = reinterpret_cast<double *>(pix);             // 64 doubles; the pixel bytes are dead
            const double tx_lo = D.termx[lane], tx_hi = D.termx[lane + 32];
            const uint32_t below = (1u << lane) - 1u;
            uint32_t lanes = __ballot_sync(0xffffffffu, (np_a | np_b) != 0);
            while (lanes) {
                const int l = __ffs(lanes) - 1;
                lanes &= lanes - 1;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const u64 w = shfl_u64(half ? np_b : np_a, l);
                    if (!w) continue;                                    // warp-uniform
                    const uint32_t wl = (uint32_t)w, wh = (uint32_t)(w >> 32);
                    const int n_lo = __popc(wl), n = n_lo + __popc(wh);
                    __syncwarp();                                       // previous row's reads are done
                    if ((wl >> lane) & 1) tbuf[__popc(wl & below)] = tx_lo;
                    if ((wh >> lane) & 1) tbuf[n_lo + __popc(wh & below)] = tx_hi;
                    __syncwarp();
                    const double ty = D.termy[2 * l + half];
MELF_UNROLL(MELF_DIALS_U4)
                    for (int i = 0; i < n; ++i) {
                        mom_x = __dadd_rn(mom_x, tbuf[i]);
                        mom_y = __dadd_rn(mom_y, ty);
                    }
                }
            }
            __syncwarp();
        }
        const double sgn = D.neg_mom ? -1.0 : 1.0;
        mom_angle = angle_by_vector(sgn * mom_x, sgn * mom_y);

        if (tl) tl[5] = clock64();
        // ---- ring pixels (_reading.py:51-96) ------------------------------------------
        // One traversal of the (angle, squared distance)-sorted ring table compacts the indices
        // of the kept entries into shared memory; the first kept entry has the minimum angle.
        nm_s[wslot][2 * lane] = nm_a;
        nm_s[wslot][2 * lane + 1] = nm_b;
        __syncwarp();
        const uint16_t *rpos = A.ring_pos + D.ring_off;
        const double *rang = A.ring_angle + D.ring_off;
        const double *rsqd = A.ring_sqd + D.ring_off;
        uint16_t *klist = reinterpret_cast<uint16_t *>(pix + 512);     // <= n_ring entries (host sizes pix_bytes)
        double min_angle = 0.0;
        int n_unwrapped = 0;
        int rpos_nx = lane < D.n_ring ? (int)rpos[lane] : 0;      // fetched one iteration ahead
MELF_UNROLL(MELF_DIALS_U2)
        for (int base = 0; base < D.n_ring; base += 32) {
            const int e = base + lane;
            bool kept = false, wrapped = false;
            double a = 0.0;
            bool inmask = false;
            const int pos = rpos_nx;
            rpos_nx = e + 32 < D.n_ring ? (int)rpos[e + 32] : 0;
            if (e < D.n_ring) inmask = (nm_s[wslot][pos >> 8] >> (pos & 255)) & 1;
            // the table is sorted by angle and the needle crosses the ring in a narrow sector: most
            // groups of 32 entries have no pixel in the needle mask at all
            if (!__any_sync(0xffffffffu, inmask)) continue;
            if (inmask) {
                a = rang[e];
                kept = near_momentum(a, mom_angle);
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, kept);
            if (n_kept == 0 && bal) min_angle = __shfl_sync(0xffffffffu, a, __ffs(bal) - 1);
            if (kept) {
                wrapped = !(fabs(__dsub_rn(a, min_angle)) < 0.75);
                klist[n_kept + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)e;
            }
            // wrapped entries have the largest angles, i.e. they form the tail of the kept list
            n_unwrapped += __popc(__ballot_sync(0xffffffffu, kept && !wrapped));
            n_kept += __popc(bal);
        }
        __syncwarp();
        if (n_kept == 0) {
            res.flag = 2;                                // _reading.py:79-81
        } else {
            double num = 0.0, den = 0.0;
            if (n_kept >= 5) {
                // sorted()[cut:-cut] (_reading.py:86-91): the wrapped entries ("a - 1", negative
                // keys) come first, then the others, each group in table order
                const int cut = min(2, (n_kept - 3) / 2);
                const int n_wrapped = n_kept - n_unwrapped;
MELF_UNROLL(MELF_DIALS_U4)
                for (int k = cut; k < n_kept - cut; ++k) {
                    const bool w = k < n_wrapped;
                    const int e = klist[w ? n_unwrapped + k : k - n_wrapped];
                    double aa = rang[e];
                    if (w) aa = __dsub_rn(aa, 1.0);
                    const double dd = rsqd[e];
                    num = __dadd_rn(num, __dmul_rn(aa, dd));
                    den = __dadd_rn(den, dd);
                }
            } else {
                // fewer than 5: no sort, the reference sums in raster order of the points
                double ea[4], ed[4];
                int ep[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    ea[q] = 0.0; ed[q] = 0.0; ep[q] = 0x7fffffff;
                    if (q < n_kept) {
                        const int e = klist[q];
                        double aa = rang[e];
                        if (!(fabs(__dsub_rn(aa, min_angle)) < 0.75)) aa = __dsub_rn(aa, 1.0);
                        ea[q] = aa; ed[q] = rsqd[e]; ep[q] = rpos[e];
                    }
                }
                // insertion sort of <= 4 entries by raster position
#pragma unroll
                for (int i = 1; i < 4; ++i)
#pragma unroll
                    for (int j = i; j > 0; --j)
                        if (j < n_kept && ep[j] < ep[j - 1]) {
                            const int tp = ep[j]; ep[j] = ep[j - 1]; ep[j - 1] = tp;
                            const double ta = ea[j]; ea[j] = ea[j - 1]; ea[j - 1] = ta;
                            const double td = ed[j]; ed[j] = ed[j - 1]; ed[j - 1] = td;
                        }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q < n_kept) {
                        num = __dadd_rn(num, __dmul_rn(ea[q], ed[q]));
                        den = __dadd_rn(den, ed[q]);
                    }
            }
            const double angle = __ddiv_rn(num, den);
            const double fixed = __dsub_rn(angle, D.aoz_turns);
            res.pos = py_mod_pos(__dmul_rn(10.0, fixed), 10.0);
        }
    }

    if (tl) tl[6] = clock64();
    // ---- taps --------------------------------------------------------------------------
    const size_t fd = (size_t)f * A.n_dials + d;
    if (A.taps.d_mask_orig) { A.taps.d_mask_orig[fd * 64 + 2 * lane] = m0a; A.taps.d_mask_orig[fd * 64 + 2 * lane + 1] = m0b; }
    if (A.taps.d_mask_de) { A.taps.d_mask_de[fd * 64 + 2 * lane] = era; A.taps.d_mask_de[fd * 64 + 2 * lane + 1] = erb; }
    if (A.taps.d_needle_mask) { A.taps.d_needle_mask[fd * 64 + 2 * lane] = nm_a; A.taps.d_needle_mask[fd * 64 + 2 * lane + 1] = nm_b; }
    if (lane == 0) {
        if (A.taps.d_dial_color) {
            A.taps.d_dial_color[fd * 3 + 0] = col_h; A.taps.d_dial_color[fd * 3 + 1] = col_l; A.taps.d_dial_color[fd * 3 + 2] = col_s;
        }
        if (A.taps.d_counts) {
            A.taps.d_counts[fd * 4 + 0] = n_needle; A.taps.d_counts[fd * 4 + 1] = n_kept;
            A.taps.d_counts[fd * 4 + 2] = best_area2; A.taps.d_counts[fd * 4 + 3] = ncomp;
        }
        if (A.taps.d_momentum) {
            A.taps.d_momentum[fd * 3 + 0] = mom_x; A.taps.d_momentum[fd * 3 + 1] = mom_y; A.taps.d_momentum[fd * 3 + 2] = mom_angle;
        }
    }

    // ---- frame result (_reading.py:98-115, 163-182): the last dial warp of the frame ------
    if (lane == 0) {
        MelfDialSlot *S = A.slots + f;
        S->pos[d] = res.pos;
        S->flag[d] = res.flag;
        __threadfence();
        const uint32_t prev = atomicAdd(&S->count, 1u);
        if (prev != (uint32_t)A.n_dials - 1u) return;
        __threadfence();
        S->count = 0;                                   // ready for the next launch
        DialOut dial_out[MELF_MAX_DIALS];
        for (int i = 0; i < A.n_dials; i++) {
            dial_out[i].pos = __ldcg(&S->pos[i]);
            dial_out[i].flag = __ldcg(&S->flag[i]);
        }
        const double nan = __longlong_as_double(0x7ff8000000000000ll);
        melf_result r;
        r.match_x = mx; r.match_y = my; r.max_val = max_val;
        r.failed_dial = -1; r.unreadable_mask = 0; r.value = nan;
        for (int i = 0; i < MELF_MAX_DIALS; i++) r.dial_pos[i] = nan;
        int first_fail = -1;
        for (int i = 0; i < A.n_dials; i++)
            if (dial_out[i].flag == 1) { first_fail = i; break; }
        const int upto = first_fail >= 0 ? first_fail : A.n_dials;
        for (int i = 0; i < upto; i++) {
            if (dial_out[i].flag == 2) r.unreadable_mask |= 1u << i;
            else r.dial_pos[i] = dial_out[i].pos;
        }
        if (first_fail >= 0) {
            r.status = MELF_ST_NO_CONTOURS;              // raised at the first such dial
            r.failed_dial = first_fail;
            r.unreadable_mask = 0;
        } else if (r.unreadable_mask) {
            r.status = MELF_ST_ANGLE_UNDETERMINED;
        } else {
            r.status = MELF_ST_OK;
            if (A.n_dials == 4) {
                const double r4 = r.dial_pos[A.value_order[0]], r3 = r.dial_pos[A.value_order[1]];
                const double r2 = r.dial_pos[A.value_order[2]], r1 = r.dial_pos[A.value_order[3]];
                const int d3 = carry_digit(r3, r4);
                const int d2 = carry_digit(r2, (double)d3);
                const int d1 = carry_digit(r1, (double)d2);
                r.value = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(d1, 100.0), __dmul_rn(d2, 10.0)),
                                              __dmul_rn(d3, 1.0)), __ddiv_rn(r4, 10.0));
            }
        }
        A.out[f] = r;
    }
}

// host launcher (called from melf_api.cu)
cudaError_t melf_launch_dials(const MelfDialsArgs &args, int n_frames, cudaStream_t stream)
{
    const int smem = DIALS_WPB * args.pix_bytes;
    cudaError_t e = cudaFuncSetAttribute(melf_dials_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         smem > 48 * 1024 ? smem : 48 * 1024);
    if (e != cudaSuccess) return e;
    const unsigned grid = (unsigned)(((n_frames + DIALS_WPB - 1) / DIALS_WPB) * args.n_dials);
    melf_dials_kernel<<<grid, 32 * DIALS_WPB, smem, stream>>>(args);
    return cudaGetLastError();
}
