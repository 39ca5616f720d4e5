"""Pure-Python model of the per-dial CUDA kernel (``csrc/melf_dials.cu``).

TEST INFRASTRUCTURE.  It mirrors the kernel's *formulation* -- 64-bit row words,
bit-parallel run filling, 8-connected flood for components, 4-connected exterior flood
for hole filling, bit-quad area, statically sorted ring table with a wrapped/unwrapped
partition, sequential float64 sums -- so that this formulation can be checked against
the oracle (Suzuki border following + shoelace + polygon fill + qsort) on the CPU, with
no GPU.  The static tables come from the product library's host code
(``melf_host_dial_tables``), so that code is covered too.
"""
import math

M64 = (1 << 64) - 1


def brev64(x):
    return int('{:064b}'.format(x)[::-1], 2)


def spread(seed, m):
    """Fill every run of ones of m that contains a seed bit (both directions)."""
    s = seed & m
    up = ((((m + s) & M64) ^ m) | s) & m
    mr, sr = brev64(m), brev64(s)
    dn = brev64(((((mr + sr) & M64) ^ mr) | sr) & mr)
    return up | dn


def h3(x):
    return (x | (x << 1) | (x >> 1)) & M64


def e3(x):
    return x & ((x << 1) | 1) & ((x >> 1) | (1 << 63)) & M64


def flood8(seed_rows, allowed):
    cur = list(seed_rows)
    while True:
        new = [spread(cur[r], allowed[r]) for r in range(64)]
        nxt = list(new)
        for r in range(64):
            up = new[r - 1] if r > 0 else 0
            dn = new[r + 1] if r < 63 else 0
            nxt[r] |= h3(up | dn) & allowed[r]
        if nxt == cur:
            return cur
        cur = nxt


def fill_holes(comp):
    """comp plus every background region not 4-connected to the outside of its bbox."""
    rows = [r for r in range(64) if comp[r]]
    rmin, rmax = rows[0], rows[-1]
    union = 0
    for w in comp:
        union |= w
    cmin = (union & -union).bit_length() - 1
    cmax = union.bit_length() - 1
    colrange = ((1 << (cmax + 1)) - 1) & ~((1 << cmin) - 1)
    allowed = [~w & M64 for w in comp]
    reach = [(~colrange & M64) if rmin <= r <= rmax else M64 for r in range(64)]
    while True:
        new = [reach[r] | spread(reach[r] & allowed[r], allowed[r]) for r in range(64)]
        nxt = list(new)
        for r in range(64):
            up = new[r - 1] if r > 0 else M64
            dn = new[r + 1] if r < 63 else M64
            nxt[r] |= (up | dn) & allowed[r]
        if nxt == reach:
            break
        reach = nxt
    return [~w & M64 for w in reach]


def area2(filled):
    total = 0
    for r in range(-1, 64):
        x = filled[r] if r >= 0 else 0
        y = filled[r + 1] if r < 63 else 0
        q0, q1, q2, q3 = x, x >> 1, y, y >> 1
        n4 = bin(q0 & q1 & q2 & q3).count('1')
        t01, x01, t23, x23 = q0 & q1, q0 ^ q1, q2 & q3, q2 ^ q3
        n3 = bin((t01 & x23) | (x01 & t23)).count('1')
        total += 2 * n4 + n3
    return total


def py_mod(a, m):
    r = math.fmod(a, m)
    if r != 0.0:
        if r < 0.0:
            r += m
    else:
        r = math.copysign(0.0, m)
    return r


def angle_by_vector(x, y):
    if y == 0:
        return 0.25 if x > 0 else 0.75 if x < 0 else None
    at = math.atan(x / y) / (2 * math.pi)
    return py_mod(-at + (0.5 if y > 0 else 0.0), 1.0)


def read_dial(T, m0, neg_mom, aoz_turns):
    """T: dict of static tables for one dial (see melf_host_dial_tables);
    m0: 64 row words of the inRange mask over the window (bits outside the rect 0).
    Returns dict(status, pos, needle_mask rows, mask_de rows, n_needle, n_kept, area2)."""
    V, RV = T['colvalid'], T['rowvalid']

    def rv(r):
        return (RV >> r) & 1 if 0 <= r < 64 else 0

    hh = [h3(w) for w in m0]
    dil = [0] * 64
    for r in range(64):
        up = hh[r - 1] if r > 0 else 0
        dn = hh[r + 1] if r < 63 else 0
        dil[r] = (up | hh[r] | dn) & V if rv(r) else 0
    pe = [e3(dil[r] | (~V & M64)) if rv(r) else M64 for r in range(64)]
    ero = [0] * 64
    for r in range(64):
        up = pe[r - 1] if r > 0 else M64
        dn = pe[r + 1] if r < 63 else M64
        ero[r] = (up & pe[r] & dn) & V if rv(r) else 0

    remaining = [ero[r] & T['disc'][r] for r in range(64)]
    best, best_a2, ncomp = None, -1, 0
    while any(remaining):
        r0 = next(r for r in range(64) if remaining[r])
        seed = [0] * 64
        seed[r0] = remaining[r0] & -remaining[r0]
        comp = flood8(seed, remaining)
        filled = fill_holes(comp)
        a2 = area2(filled)
        ncomp += 1
        if a2 > best_a2:
            best, best_a2 = filled, a2
        remaining = [remaining[r] & ~filled[r] & M64 for r in range(64)]
    out = dict(mask_de=ero, ncomp=ncomp, area2=best_a2, pos=None)
    if best is None:
        out['status'] = 'no_contours'
        return out
    nm = best if best_a2 > 200 else ero
    out['needle_mask'] = nm

    mx = my = 0.0
    npts = 0
    for r in range(64):
        w = nm[r] & T['disc'][r]
        while w:
            c = (w & -w).bit_length() - 1
            w &= w - 1
            mx += T['termx'][c]
            my += T['termy'][r]
            npts += 1
    out['n_needle'] = npts
    sgn = -1 if neg_mom else 1
    mom = angle_by_vector(sgn * mx, sgn * my)
    out['momentum'] = (mx, my, mom)

    kept = []
    if mom is not None:
        for (pos, a, d) in zip(T['ring_pos'], T['ring_angle'], T['ring_sqd']):
            (row, col) = (pos >> 8, pos & 255)
            if not (nm[row] >> col) & 1:
                continue
            diff = abs(a - mom)
            if min(diff, abs(diff - 1)) < 0.25:
                kept.append((pos, a, d))
    n = len(kept)
    out['n_kept'] = n
    if n == 0:
        out['status'] = 'unreadable'
        return out
    min_angle = kept[0][1]          # the table is sorted by angle
    wrapped = [(p, a - 1, d) for (p, a, d) in kept if not abs(a - min_angle) < 0.75]
    straight = [(p, a, d) for (p, a, d) in kept if abs(a - min_angle) < 0.75]
    if n >= 5:
        cut = min(2, (n - 3) // 2)
        order = (wrapped + straight)[cut:n - cut]
    else:
        order = sorted(wrapped + straight)      # raster order (pos = row<<8 | col)
    num = den = 0.0
    for (_p, a, d) in order:
        num = num + a * d
        den = den + d
    angle = num / den
    out['pos'] = py_mod(10.0 * (angle - aoz_turns), 10.0)
    out['status'] = 'ok'
    return out


# ---------------------------------------------------------------------------------------
# Model of the tensor-core TM_CCOEFF kernel (csrc/melf_match_tc.cu): the same shared-memory
# byte layouts (16-byte column strips), the same per-step operand addressing (a row shift
# of the lightness slab = a 16-byte step of the descriptor start address), the same
# Toeplitz operand image (built by the product library's host code; A operand, row
# m = 2 n + q), the slab as the N-side operand with N = round16(h_c + 1), and the
# D[2x][y] + D[2x+1][y+1] epilogue.  numpy int64 stands in for tcgen05.mma.
# ---------------------------------------------------------------------------------------
SLAB_ROWS = 264


def match_tc_model(light, toeplitz, tmpl_w, tmpl_h, plan):
    """light: (crop_h, crop_w) u8; toeplitz: bytes from melf_host_match_tc_toeplitz;
    plan = (n_tx, n_ty, chunk_h, n_pairs).  Returns sum T*L for every result position."""
    import numpy as np
    (n_tx, n_ty, chunk_h, n_pairs) = plan
    (crop_h, crop_w) = light.shape
    (res_h, res_w) = (crop_h - tmpl_h + 1, crop_w - tmpl_w + 1)
    toep = np.frombuffer(toeplitz, np.uint8).reshape(n_pairs, 16, 128, 16).astype(np.int64)
    out = np.full((res_h, res_w), -1, np.int64)
    for ty in range(n_ty):
        for tx in range(n_tx):
            (x0, y0) = (tx * 64, ty * chunk_h)
            (w_c, h_c) = (min(64, res_w - x0), min(chunk_h, res_h - y0))
            assert 1 <= h_c <= 137 and h_c + tmpl_h - 1 <= 255
            n_cols = (h_c + 1 + 15) // 16 * 16
            # slab[strip][row][16]: zero outside the crop
            slab = np.zeros((16, SLAB_ROWS, 16), np.int64)
            for s in range(16):
                for b in range(16):
                    cx = x0 + 16 * s + b
                    if cx < crop_w:
                        rows = min(SLAB_ROWS, crop_h - y0)
                        slab[s, :rows, b] = light[y0:y0 + rows, cx]
            acc = np.zeros((128, n_cols), np.int64)
            for p in range(n_pairs):
                row0 = 2 * p
                assert row0 + n_cols <= SLAB_ROWS
                for ks in range(8):
                    # A[m][k] = toep[p][2ks + k // 16][m][k % 16]; B[r][k] = slab[2ks + k // 16][row0 + r][k % 16]
                    a = np.concatenate([toep[p, 2 * ks], toep[p, 2 * ks + 1]], axis=1)
                    b = np.concatenate([slab[2 * ks, row0:row0 + n_cols], slab[2 * ks + 1, row0:row0 + n_cols]], axis=1)
                    acc += a @ b.T
            for yl in range(h_c):
                out[y0 + yl, x0:x0 + w_c] = acc[0:2 * w_c:2, yl] + acc[1:2 * w_c:2, yl + 1]
    assert (out >= 0).all()
    return out
