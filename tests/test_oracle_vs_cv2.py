"""Every primitive of the CPU oracle against the installed cv2 (IPP off) on random inputs.

The oracle restates the published algorithms of the OpenCV calls on the path
(oracle/melf_oracle.c); the reference pins opencv-python 3.4.5.20, this image has 4.13 --
both pad the image before border following (copyMakeBorder, since 3.2), so components that
touch the image frame are contours like any other.  That case is covered here explicitly:
the sample discs stay >= 7 px away from the rect's frame and no golden frame can see it.
"""
import numpy as np
import pytest

import oracle

cv2 = pytest.importorskip('cv2')
cv2.ipp.setUseIPP(False)
cv2.setNumThreads(1)


def _cv2_select(mask_de, dial_mask):
    """_reading.py:132-148 with cv2 itself."""
    res = cv2.findContours(mask_de & dial_mask, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_NONE)
    contours = res[-2]
    if not len(contours):
        return (False, None, -1.0)
    contour = sorted(contours, key=cv2.contourArea)[-1]
    area = cv2.contourArea(contour)
    if area > 100:
        nm = np.zeros_like(mask_de)
        cv2.drawContours(nm, [contour], -1, 255, -1)
    else:
        nm = mask_de
    return (True, nm, area)


def _check_select(mask_de, dial_mask, context):
    (have_c, nm_c, area_c) = _cv2_select(mask_de, dial_mask)
    (have_o, nm_o, area_o) = oracle.select_needle_mask(mask_de, dial_mask)
    assert have_o == have_c, context
    if have_c:
        assert area_o == area_c, (context, area_o, area_c)
        assert np.array_equal(nm_o, nm_c), context


def test_components_touching_the_image_frame():
    """a 14x14 blob on the top edge has contour area 169 (13x13), not 156."""
    full = np.full((40, 50), 255, np.uint8)
    m = np.zeros((40, 50), np.uint8)
    m[0:14, 10:24] = 255
    (have, nm, area) = oracle.select_needle_mask(m, full)
    assert have and area == 169.0 and int(np.count_nonzero(nm)) == 196
    _check_select(m, full, 'top edge blob')
    # a lone frame pixel is a contour (area 0 -> the <= 100 branch keeps the closed mask)
    for (y, x) in ((0, 0), (0, 17), (39, 49), (20, 0), (20, 49), (39, 3)):
        m = np.zeros((40, 50), np.uint8)
        m[y, x] = 255
        (have, nm, area) = oracle.select_needle_mask(m, full)
        assert have and area == 0.0
        _check_select(m, full, ('lone pixel', y, x))
    # U shape whose opening is closed only by the image frame: the notch is not a hole
    m = np.zeros((40, 50), np.uint8)
    m[0:20, 5:9] = 255
    m[0:20, 25:29] = 255
    m[16:20, 5:29] = 255
    _check_select(m, full, 'U on the top edge')
    # ring cut by the corner
    m = np.zeros((40, 50), np.uint8)
    cv2.circle(m, (3, 3), 14, 255, 3)
    _check_select(m, full, 'ring in the corner')
    # everything set
    _check_select(full.copy(), full, 'all set')


def test_select_needle_mask_random_masks_incl_frame():
    rng = np.random.default_rng(77)
    for it in range(300):
        h, w = int(rng.integers(8, 70)), int(rng.integers(8, 90))
        c = int(rng.integers(1, 7))
        coarse = rng.random((h // c + 2, w // c + 2)) < rng.uniform(0.15, 0.7)
        m = (np.kron(coarse, np.ones((c, c), bool))[:h, :w] ^ (rng.random((h, w)) < 0.04))
        mask_de = m.astype(np.uint8) * 255
        if it % 3 == 0:
            dial = np.full((h, w), 255, np.uint8)        # disc reaching every edge
        else:
            dial = np.zeros((h, w), np.uint8)
            cv2.circle(dial, (int(rng.integers(0, w)), int(rng.integers(0, h))),
                       int(rng.integers(3, max(h, w))), 255, -1)
        _check_select(mask_de, dial, it)


def test_morph3_is_cv2_dilate_erode():
    rng = np.random.default_rng(5)
    k = np.ones((3, 3), np.uint8)
    for it in range(100):
        h, w = int(rng.integers(1, 60)), int(rng.integers(1, 80))
        m = ((rng.random((h, w)) < rng.uniform(0.05, 0.9)) * 255).astype(np.uint8)
        assert np.array_equal(oracle.morph3(m, False), cv2.dilate(m, k)), it
        assert np.array_equal(oracle.morph3(m, True), cv2.erode(m, k)), it


def test_convert_to_hls_is_cv2_cvtcolor():
    rng = np.random.default_rng(6)
    for (h, w) in ((1, 1), (3, 7), (17, 8), (33, 250), (20, 641)):
        bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        for shift in (0, 128, 255):
            want = cv2.cvtColor(bgr, cv2.COLOR_BGR2HLS_FULL) + np.array([shift, 0, 0], np.uint8)
            assert np.array_equal(oracle.convert_to_hls(bgr, shift), want), (h, w, shift)
    # greys, saturated primaries, extremes
    edge = np.array([[[v, v, v] for v in (0, 1, 127, 128, 254, 255)] +
                     [[255, 0, 0], [0, 255, 0], [0, 0, 255], [255, 255, 0], [0, 255, 255], [255, 0, 255],
                      [1, 0, 0], [0, 0, 1], [254, 255, 255], [255, 254, 0]]], np.uint8)
    edge = np.repeat(edge, 3, axis=0)
    assert np.array_equal(oracle.convert_to_hls(edge, 0), cv2.cvtColor(edge, cv2.COLOR_BGR2HLS_FULL))


def test_match_template_against_cv2_within_tolerance():
    """Exact integer TM_CCOEFF vs cv2's float32 DFT: the scores agree within a stated tolerance
    (DESIGN.md section 2); the arg-max agrees whenever the best exact score leads the runner-up by
    more than twice cv2's error on that map."""
    rng = np.random.default_rng(8)
    for (ih, iw, th, tw) in ((40, 60, 9, 13), (120, 90, 31, 22), (250, 250, 119, 188)):
        img = rng.integers(0, 256, (ih, iw), dtype=np.uint8)
        tmpl = rng.integers(0, 256, (th, tw), dtype=np.uint8)
        (y0, x0) = (int(rng.integers(0, ih - th + 1)), int(rng.integers(0, iw - tw + 1)))
        img[y0:y0 + th, x0:x0 + tw] = (tmpl.astype(np.int16) // 2 + 60).astype(np.uint8)
        (mx, my, mv, res) = oracle.match_template(img, tmpl)
        ref = cv2.matchTemplate(img, tmpl, cv2.TM_CCOEFF)
        err = float(np.max(np.abs(ref.astype(np.float64) - res.astype(np.float64))))
        scale = float(np.max(np.abs(res)))
        assert err <= max(4.0, 2e-5 * scale), (err, scale)
        (_mn, cmax, _mnl, cloc) = cv2.minMaxLoc(ref)
        assert (mx, my) == (x0, y0) == tuple(cloc)
        assert abs(float(mv) - cmax) <= err + 1e-3


def test_mean_round_and_inrange_semantics():
    """get_dial_color (cv2.mean + round) and inRange inclusive bounds, via read_dials_hls taps."""
    import sys
    import os
    sys.path.insert(0, os.path.dirname(__file__))
    from helpers import golden_set
    gs = golden_set(1)
    P = gs.oracle_params(threshold=-1e30)
    rng = np.random.default_rng(9)
    hls = rng.integers(0, 256, (gs.tmpl_h, gs.tmpl_w, 3), dtype=np.uint8)
    (_res, taps) = oracle.read_dials_hls(P, hls, want_taps=True)
    for (i, d) in enumerate(gs.dials):
        (cx, cy) = (int(d['center'][0]), int(d['center'][1]))
        mean = cv2.mean(hls[cy - 2:cy + 3, cx - 2:cx + 3])[:3]
        col = [int(round(v)) for v in mean]
        assert list(taps.dial_color[i]) == col
        lo = np.array([max(c - r, 0) for (c, r) in zip(col, d['color_range'])], np.uint8)
        hi = np.array([min(c + r, 255) for (c, r) in zip(col, d['color_range'])], np.uint8)
        assert np.array_equal(taps.mask_orig[i], cv2.inRange(hls, lo, hi))
