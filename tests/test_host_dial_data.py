"""``meterelf_b200._dial_data`` against cv2 and against the reference's masks.

The reference builds the per-dial disc and annulus masks with ``cv2.circle`` + ``cv2.floodFill``
(``_dial_data.py:22-55``); the product restates both in numpy.  Here: the drawing primitives
against cv2 for radii 1..80 (clipped circles included), the flood fill against ``cv2.floodFill``,
and ``get_dial_data`` for both sample params against the masks the unmodified reference produced
(``ref_mask`` / ``ref_circle_mask`` in ``tests/golden/set*_params.npz``, written by
``tests/golden/make_golden.py``).
"""
import numpy as np
import pytest

from helpers import golden_set, write_params_file
from meterelf_b200 import _dial_data, _params


def test_get_dial_data_equals_the_reference_masks(tmp_path):
    for set_no in (1, 2):
        gs = golden_set(set_no)
        d = tmp_path / ('set%d' % set_no)
        d.mkdir()
        params = _params.load(write_params_file(gs, d))
        data = _dial_data.get_dial_data(params)
        assert list(sorted(data)) == sorted(gs.names)
        for ref in gs.dials:
            dd = data[ref['name']]
            assert dd.name == ref['name'] and tuple(dd.center) == tuple(ref['center'])
            assert dd.mask.dtype == np.uint8 and dd.mask.shape == (gs.tmpl_h, gs.tmpl_w)
            assert np.array_equal(dd.mask, ref['mask']), (set_no, ref['name'], 'mask')
            assert np.array_equal(dd.circle_mask, ref['circle_mask']), (set_no, ref['name'], 'circle_mask')
        # cached per Params object like the reference (_dial_data.py:14-19)
        assert _dial_data.get_dial_data(params) is data


def test_draw_circle_is_cv2_circle_for_all_radii_incl_clipped():
    cv2 = pytest.importorskip('cv2')
    rng = np.random.default_rng(3)
    (h, w) = (90, 120)
    centres = [(60, 45), (0, 0), (119, 89), (3, 80), (115, 2), (-5, 40), (60, -7), (130, 95)]
    for radius in range(0, 81):
        for c in centres + [(int(rng.integers(-10, w + 10)), int(rng.integers(-10, h + 10)))]:
            got = np.zeros((h, w), np.uint8)
            want = np.zeros((h, w), np.uint8)
            _dial_data.draw_circle(got, c, radius)
            cv2.circle(want, c, radius, 255)
            assert np.array_equal(got, want), (radius, c)


def test_flood_fill_is_cv2_floodfill():
    cv2 = pytest.importorskip('cv2')
    rng = np.random.default_rng(4)
    for it in range(60):
        (h, w) = (int(rng.integers(5, 60)), int(rng.integers(5, 80)))
        img = ((rng.random((h, w)) < rng.uniform(0.2, 0.6)) * 255).astype(np.uint8)
        if it % 2:                      # ring shapes like the dial outlines
            img[:] = 0
            cv2.circle(img, (w // 2, h // 2), min(h, w) // 3, 255)
            cv2.circle(img, (w // 2, h // 2), min(h, w) // 5, 255)
        seed = (int(rng.integers(0, w)), int(rng.integers(0, h)))
        got = img.copy()
        want = img.copy()
        _dial_data.flood_fill(got, seed)
        cv2.floodFill(want, None, seed, 255)
        assert np.array_equal(got, want), (it, seed)
    with pytest.raises(ValueError):
        _dial_data.flood_fill(np.zeros((4, 4), np.uint8), (4, 0))


def test_dial_masks_equal_the_reference_recipe_for_random_geometry():
    """Random template sizes, centres (also near / on the frame), radii and ring thicknesses
    (1 and 2 included: the outlines coincide or touch): ``_dial_data.py:22-55`` restated with cv2
    -- shared floodFill mask and all -- against the numpy recipe of the product."""
    cv2 = pytest.importorskip('cv2')

    def with_cv2(th, tw, center, start_radius, thickness):
        mask = np.zeros((th, tw), np.uint8)
        for i in [0, thickness - 1]:
            cv2.circle(mask, center, start_radius + i, 255)
        fill_mask = np.zeros((th + 2, tw + 2), np.uint8)
        cv2.floodFill(mask, fill_mask, (center[0] + start_radius + 1, center[1]), 255)
        circle_mask = mask.copy()
        cv2.floodFill(mask, fill_mask, center, 255)
        return (mask, circle_mask)

    def with_numpy(th, tw, center, inner_radius, thickness):
        canvas = np.zeros((th, tw), np.uint8)
        for extra in (0, thickness - 1):
            _dial_data.draw_circle(canvas, center, inner_radius + extra)
        _dial_data.flood_fill(canvas, (center[0] + inner_radius + 1, center[1]))
        circle_mask = canvas.copy()
        _dial_data.flood_fill(canvas, center)
        return (canvas, circle_mask)

    rng = np.random.default_rng(1)
    for it in range(400):
        (th, tw) = (int(rng.integers(30, 130)), int(rng.integers(30, 200)))
        c = (int(rng.integers(0, tw)), int(rng.integers(0, th)))
        (r, t) = (int(rng.integers(1, 40)), int(rng.integers(1, 10)))
        try:
            want = with_cv2(th, tw, c, r, t)
        except cv2.error:                        # seed outside of the image
            with pytest.raises(ValueError):
                with_numpy(th, tw, c, r, t)
            continue
        got = with_numpy(th, tw, c, r, t)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (th, tw, c, r, t)
