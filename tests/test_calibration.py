"""Offline calibration: the average meter image and the dial centres against golden
vectors produced by the unmodified reference (tests/golden/make_golden_calibration.py)."""
import os

import numpy as np
import pytest

import oracle
from helpers import GOLDEN, golden_set, write_params_file

# the reference's own expectations, tests/test_meterelf.py:140-145
EXPECTED_CENTER_DATA = [(37.4, 63.5, 14), (94.5, 86.3, 15), (135.6, 71.5, 13), (161.0, 36.5, 13)]
FULL = os.path.join(GOLDEN, '_full', 'set1_crops.npz')


def _golden():
    return np.load(os.path.join(GOLDEN, 'set1_calibration.npz'))


def _subset_crops():
    G = _golden()
    return golden_set(1).subset()['crops'][G['subset_index']]


def _oracle_match_xy(gs, crops):
    xy = []
    for crop in crops:
        light = np.ascontiguousarray(oracle.convert_to_hls(crop, gs.hue_shift)[..., 1])
        (x, y, _val, _map) = oracle.match_template(light, gs.template, want_map=False)
        xy.append((x, y))
    return np.array(xy, dtype=np.int32)


def test_golden_reproduces_the_reference_tests_expectations():
    G = _golden()
    assert len(G['full_names']) == 79
    for ((cx, cy), d, (ex, ey, ed)) in zip(G['full_centers'], G['full_diameters'], EXPECTED_CENTER_DATA):
        assert int(d) == ed and abs(cx - ex) < 0.05 and abs(cy - ey) < 0.05


def test_oracle_average_matches_reference_on_subset():
    gs = golden_set(1)
    G = _golden()
    crops = _subset_crops()
    avg = oracle.average_translated(crops, _oracle_match_xy(gs, crops))
    assert np.array_equal(avg, G['subset_avg'])


@pytest.mark.skipif(not os.path.exists(FULL), reason='full crop set not generated')
def test_oracle_average_matches_reference_on_all_files():
    G = _golden()
    full = np.load(FULL)
    names = [str(s) for s in full['names']]
    crops = np.stack([full['crops'][names.index(str(fn))] for fn in G['full_names']])
    avg = oracle.average_translated(crops, G['full_match_xy'])
    assert np.array_equal(avg, G['full_avg'])


def test_outer_borders_and_fit_ellipse_are_cv2s():
    """The cv2-free contour following and conic fit of the product's calibration
    (``meterelf_b200/_calibration.py``) against ``cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_NONE)``
    point for point (nested components dropped, frame pixels included) and ``cv2.fitEllipse``."""
    cv2 = pytest.importorskip('cv2')
    from meterelf_b200 import _calibration as cal
    rng = np.random.default_rng(0)
    n_fits = 0
    for it in range(160):
        (h, w) = (int(rng.integers(25, 80)), int(rng.integers(25, 100)))
        c = int(rng.integers(1, 6))
        m = np.kron(rng.random((h // c + 2, w // c + 2)) < rng.uniform(0.2, 0.7), np.ones((c, c), bool))[:h, :w]
        m = m.astype(np.uint8) * 255
        blobs = it % 2 == 0
        if blobs:                                     # hub-like shapes: filled ellipses, some overlapping
            m[:] = 0
            for _ in range(3):
                cv2.ellipse(m, (int(rng.integers(10, w)), int(rng.integers(10, h))),
                            (int(rng.integers(3, 12)), int(rng.integers(3, 12))), float(rng.uniform(0, 180)),
                            0, 360, 255, -1)
        want = cv2.findContours(m, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_NONE)[-2]
        got = cal.outer_borders(m)
        assert len(got) == len(want), it
        for (g, r) in zip(got, want):
            r = r.reshape(-1, 2)
            assert np.array_equal(g, r), it
            if blobs and len(r) >= 12:
                (rc, rs, _ra) = cv2.fitEllipse(r)
                if min(rs) < 4:
                    continue
                (gc, gsz, _ga) = cal.fit_ellipse(g)
                n_fits += 1
                assert np.allclose(gc, rc, atol=1e-4) and np.allclose(gsz, rs, atol=1e-4), (it, gc, rc, gsz, rs)
    assert n_fits > 100
    with pytest.raises(ValueError):
        cal.fit_ellipse(np.zeros((4, 2)))


def _params(tmp_path):
    from meterelf_b200 import _params
    return _params.load(write_params_file(golden_set(1), tmp_path))


@pytest.mark.gpu
def test_gpu_average_and_centers_match_reference_on_subset(tmp_path):
    from meterelf_b200 import _calibration
    params = _params(tmp_path)
    G = _golden()
    avg = _calibration.average_of_crops(params, list(_subset_crops()))
    assert avg.dtype == np.uint8 and np.array_equal(avg, G['subset_avg'])
    centers = _calibration.find_dial_centers_from_image(params, avg)
    assert [c.diameter for c in centers] == [int(d) for d in G['subset_diameters']]
    # the product's own conic fit on the product's own contours: float32 results like cv2.fitEllipse's
    assert np.allclose(np.array([c.center for c in centers]), G['subset_centers'], atol=1e-3)
    assert centers == sorted(centers, key=lambda c: c.center[0])


@pytest.mark.gpu
def test_gpu_average_raises_for_a_frame_without_dials(tmp_path):
    from meterelf_b200 import _calibration
    from meterelf_b200.exceptions import DialsNotFoundError
    params = _params(tmp_path)
    sub = golden_set(1).subset()
    names = [str(s) for s in sub['names']]
    bad = names.index('20180814021309-01-e01.jpg')
    with pytest.raises(DialsNotFoundError):
        _calibration.average_of_crops(params, [sub['crops'][0], sub['crops'][bad]], ['a.jpg', 'bad.jpg'])


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(FULL), reason='full crop set not generated')
def test_gpu_find_dial_centers_like_the_reference_test(tmp_path):
    """tests/test_meterelf.py:118-137 with the decoded crops of the 79 usable files."""
    from meterelf_b200 import _calibration
    params = _params(tmp_path)
    G = _golden()
    full = np.load(FULL)
    names = [str(s) for s in full['names']]
    crops = [full['crops'][names.index(str(fn))] for fn in G['full_names']]
    avg = _calibration.average_of_crops(params, crops)
    assert np.array_equal(avg, G['full_avg'])
    result = _calibration.find_dial_centers_from_image(params, avg)
    assert len(result) == 4
    for (center_data, (ex, ey, ed)) in zip(result, EXPECTED_CENTER_DATA):
        assert center_data.diameter == ed
        assert abs(center_data.center[0] - ex) < 0.05 and abs(center_data.center[1] - ey) < 0.05
