"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle.

Bit-exact for everything integer/byte/index: HLS planes, lightness, colour masks, match
location, per-dial masks and point counts.  The TM_CCOEFF map and max_val are compared
bit-exact as float32 too (both sides evaluate the same exact integer expression).  Dial
positions and values are float64 and compared bit-exact as well: both sides add the same
table values in the same order; the only non-shared operation is one atan per dial for
the coarse momentum direction (CUDA libm vs glibc), which feeds a < 0.25 comparison and
never the result itself.
"""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from helpers import (
    GOLDEN, SYNTHETIC_CASES, assert_results_equal, check_against_reference_record, format_result_line,
    golden_set, synthetic_batch, synthetic_expected, window_words)

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def _dev(arr):
    return _torch().from_numpy(np.ascontiguousarray(arr)).cuda()


# ---------------------------------------------------------------------------------------
# whole path, sample frames
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize('set_no', [1, 2])
def test_sample_frames_readings_identical(set_no):
    """configs[1]: readings of sample-images1/2 on the GPU diffed against the goldens."""
    gs = golden_set(set_no)
    (names, crops) = gs.frames()
    eng = gs.engine(cropped_input=True)
    got = eng.read_frames(crops)
    want = oracle.read_batch(gs.oracle_params(), crops, n_threads=8)
    n_lines = 0
    for (fn, g, w) in zip(names, got, want):
        assert_results_equal(g, w, len(gs.names), fn)
        exp = gs.expected[fn]
        line = format_result_line(g, gs.names)
        if exp['status'] == 0:
            assert line == exp['line'], fn
            n_lines += 1
        else:
            assert line.split(' (match val')[0] == exp['line'].split(' (match val')[0], fn
            assert abs(float(g['max_val']) - exp['cv2_max_val']) <= 256.0
    assert n_lines >= 19


@pytest.mark.parametrize('set_no', [1, 2])
def test_full_frames_with_meter_rect(set_no):
    """Whole 480x640 frames + meter_rect clipping on the device = host-cropped input."""
    gs = golden_set(set_no)
    sub = gs.subset()
    crops = sub['crops'][:12]
    rng = np.random.default_rng(5)
    (x0, y0, x1, y1) = gs.meter_rect
    frames = rng.integers(0, 256, size=(len(crops), 640, 480, 3), dtype=np.uint8)
    frames[:, y0:y1, x0:x1] = crops
    got = gs.engine(cropped_input=False).read_frames(frames)
    want = gs.engine(cropped_input=True).read_frames(crops)
    for (g, w) in zip(got, want):
        assert_results_equal(g, w, len(gs.names))


def test_device_resident_async_and_taps():
    torch = _torch()
    gs = golden_set(1)
    sub = gs.subset()
    names = [str(s) for s in sub['names']]
    tap_names = [str(s) for s in sub['tap_names']]
    crops = np.stack([sub['crops'][names.index(fn)] for fn in tap_names])
    (n, h, w) = crops.shape[:3]
    nd = len(gs.names)
    from meterelf_b200 import _engine
    eng = gs.engine(cropped_input=True)
    d_in = _dev(crops)
    d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device='cuda')
    t_m0 = torch.zeros((n, nd, 64), dtype=torch.int64, device='cuda')
    t_de = torch.zeros_like(t_m0)
    t_nm = torch.zeros_like(t_m0)
    t_col = torch.zeros((n, nd, 3), dtype=torch.int32, device='cuda')
    t_cnt = torch.zeros((n, nd, 4), dtype=torch.int32, device='cuda')
    t_mom = torch.zeros((n, nd, 3), dtype=torch.float64, device='cuda')
    taps = _engine.CTaps(t_m0.data_ptr(), t_de.data_ptr(), t_nm.data_ptr(), t_col.data_ptr(),
                         t_cnt.data_ptr(), t_mom.data_ptr())
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        eng.read_device(d_in.data_ptr(), n, h, w, w * 3, h * w * 3, d_out.data_ptr(),
                        stream.cuda_stream, taps)
    stream.synchronize()
    got = d_out.cpu().numpy().view(_engine.RESULT_DTYPE).reshape(n)
    O = gs.oracle_params()
    m0 = t_m0.cpu().numpy().view(np.uint64)
    de = t_de.cpu().numpy().view(np.uint64)
    nm = t_nm.cpu().numpy().view(np.uint64)
    for i in range(n):
        (res, ot) = oracle.read_frame(O, crops[i], want_taps=True)
        assert_results_equal(got[i], res, nd, tap_names[i])
        assert np.array_equal(t_col[i].cpu().numpy(), ot.dial_color)
        for d in range(nd):
            (wx, wy) = eng.dial_window(d)
            disc = window_words(gs.dials[d]['mask'], wx, wy)
            # inRange / closing are compared over the whole disc; the needle mask too
            for (mine, theirs) in ((m0, ot.mask_orig), (de, ot.mask_de), (nm, ot.needle_mask)):
                ref = window_words(theirs[d], wx, wy)
                assert np.array_equal(mine[i, d] & disc, ref & disc), (tap_names[i], d)
            cnt = t_cnt[i, d].cpu().numpy()
            assert cnt[0] == ot.n_needle_points[d]
            assert cnt[1] == ot.n_outer_kept[d]
            assert cnt[2] == int(round(2 * ot.contour_area[d]))
            mom = t_mom[i, d].cpu().numpy()
            assert mom[0] == ot.momentum[d, 0] and mom[1] == ot.momentum[d, 1]
            assert abs(mom[2] - ot.momentum[d, 2]) < 1e-15


# ---------------------------------------------------------------------------------------
# branches the sample images never reach: synthetic blob frames, permissive threshold
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize('case', [c[0] for c in SYNTHETIC_CASES])
def test_synthetic_blob_frames_all_branches(case):
    """Statuses 2 and 3, the area <= 100 branch, the tie-breaks and (set 3) discs cut by the
    rect's frame: the GPU records against the oracle's AND against what the unmodified reference
    computed on the same frames (tests/golden/synthetic_expected.json; near-tie rule for the
    few frames where cv2's float32 DFT picks another arg-max, helpers.check_against_reference_record)."""
    exp = synthetic_expected()[case]
    gs = golden_set(exp['set'])
    (x0, y0, x1, y1) = gs.meter_rect
    frames = synthetic_batch(exp['seed'], exp['n'], y1 - y0, x1 - x0)
    thr = -1e30
    got = gs.engine(cropped_input=True, threshold=thr).read_frames(frames)
    O = gs.oracle_params(threshold=thr)
    want = oracle.read_batch(O, frames, n_threads=8)
    seen = set()
    for (i, (g, wnt)) in enumerate(zip(got, want)):
        assert_results_equal(g, wnt, len(gs.names), 'frame %d' % i)
        check_against_reference_record(gs, O, frames[i], g, None, exp['frames'][i], (case, i))
        seen.add(int(wnt['status']))
    assert seen == {0, 2, 3}, seen        # the generator must reach the error branches


@pytest.mark.parametrize('set_no,seed', [(1, 303), (2, 404)])
def test_gpu_equals_the_unmodified_reference_run_live(set_no, seed, tmp_path):
    """No fixture and no port in between: the unmodified reference (``oracle/_ref``, worker processes on the host
    cores) and the CUDA path read the same freshly perturbed sample crops; status, failing / unreadable dials
    and the float64 dial positions and value agree bit for bit."""
    from helpers import write_params_file
    from test_oracle_vs_reference_live import perturbed_crops
    if not oracle.ref_available():
        pytest.skip('oracle/_ref not built')
    gs = golden_set(set_no)
    crops = perturbed_crops(gs, seed, copies=3)
    with oracle.RefPool(write_params_file(gs, tmp_path), crops, workers=4) as pool:
        (_wall, ref_records) = pool.run()
    got = gs.engine(cropped_input=True).read_frames(crops)
    assert oracle.ref_mismatches(ref_records, got, len(gs.names)) == 0
    assert sum(1 for r in ref_records if r['status'] == 0) > len(crops) // 2


def test_dial_taps_with_discs_cut_by_the_rect_frame():
    """Set 3: needle masks, point counts and contour areas of components lying on the frame of
    the dials rect (cv2.findContours pads the image) -- GPU taps against the oracle's,
    which tests/test_oracle_golden.py pins to the reference on these very frames."""
    torch = _torch()
    from meterelf_b200 import _engine
    gs = golden_set(3)
    (x0, y0, x1, y1) = gs.meter_rect
    frames = synthetic_batch(6, 64, y1 - y0, x1 - x0)
    (n, h, w) = frames.shape[:3]
    nd = len(gs.names)
    eng = gs.engine(cropped_input=True, threshold=-1e30)
    d_in = _dev(frames)
    d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device='cuda')
    t_words = [torch.zeros((n, nd, 64), dtype=torch.int64, device='cuda') for _ in range(3)]
    t_col = torch.zeros((n, nd, 3), dtype=torch.int32, device='cuda')
    t_cnt = torch.zeros((n, nd, 4), dtype=torch.int32, device='cuda')
    t_mom = torch.zeros((n, nd, 3), dtype=torch.float64, device='cuda')
    taps = _engine.CTaps(*(t.data_ptr() for t in t_words), t_col.data_ptr(), t_cnt.data_ptr(), t_mom.data_ptr())
    eng.read_device(d_in.data_ptr(), n, h, w, w * 3, h * w * 3, d_out.data_ptr(),
                    torch.cuda.current_stream().cuda_stream, taps)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy().view(_engine.RESULT_DTYPE).reshape(n)
    nm = t_words[2].cpu().numpy().view(np.uint64)
    de = t_words[1].cpu().numpy().view(np.uint64)
    cnt = t_cnt.cpu().numpy()
    O = gs.oracle_params(threshold=-1e30)
    on_frame = 0
    for i in range(n):
        (res, ot) = oracle.read_frame(O, frames[i], want_taps=True)
        assert_results_equal(got[i], res, nd, i)
        last = nd if int(res['status']) != 2 else int(res['failed_dial'])
        for d in range(last):
            (wx, wy) = eng.dial_window(d)
            disc = window_words(gs.dials[d]['mask'], wx, wy)
            for (mine, theirs) in ((de, ot.mask_de), (nm, ot.needle_mask)):
                assert np.array_equal(mine[i, d] & disc, window_words(theirs[d], wx, wy) & disc), (i, d)
            assert cnt[i, d, 0] == ot.n_needle_points[d]
            assert cnt[i, d, 2] == int(round(2 * ot.contour_area[d])), (i, d)
            m = (ot.needle_mask[d] > 0) & (gs.dials[d]['mask'] > 0)
            on_frame += int(m[0].any() or m[-1].any() or m[:, 0].any() or m[:, -1].any())
    assert on_frame >= 20, on_frame       # the case must really put needle pixels on the frame


def test_threshold_is_strict_less_than():
    gs = golden_set(2)
    crop = gs.subset()['crops'][:1]
    base = gs.engine(cropped_input=True).read_frames(crop)[0]
    assert int(base['status']) == 0
    v = float(base['max_val'])
    at = gs.engine(cropped_input=True, threshold=v).read_frames(crop)[0]
    above = gs.engine(cropped_input=True, threshold=float(np.nextafter(np.float32(v), np.float32(np.inf)))).read_frames(crop)[0]
    assert int(at['status']) == 0              # max_val < threshold is false when equal
    assert int(above['status']) == 1


def test_empty_batch_and_too_small_frame():
    from meterelf_b200 import _engine
    gs = golden_set(2)
    eng = gs.engine(cropped_input=True)
    assert len(eng.read_frames(np.zeros((0, 135, 220, 3), np.uint8))) == 0
    with pytest.raises(_engine.EngineError) as exc:
        eng.read_frames(np.zeros((1, 100, 100, 3), np.uint8))
    assert exc.value.code == _engine.ERR_TOO_SMALL


# ---------------------------------------------------------------------------------------
# stage operators
# ---------------------------------------------------------------------------------------
def test_convert_hls_full_colour_cube_bit_exact():
    """All 2^24 BGR colours: HLS, lightness plane and a colour mask, vs the oracle."""
    torch = _torch()
    from meterelf_b200 import _engine
    v = np.arange(1 << 24, dtype=np.uint32)
    cube = np.stack([v & 255, (v >> 8) & 255, v >> 16], axis=-1).astype(np.uint8).reshape(4096, 4096, 3)
    want = oracle.convert_to_hls(cube, hue_shift=128)
    lo, hi = (116, 35, 95), (134, 125, 165)
    d_in = _dev(cube)
    d_hls = torch.empty((4096, 4096, 3), dtype=torch.uint8, device='cuda')
    d_l = torch.empty((4096, 4096), dtype=torch.uint8, device='cuda')
    d_m = torch.empty((4096, 4096), dtype=torch.uint8, device='cuda')
    _engine.convert_hls_device(0, d_in.data_ptr(), 1, 4096, 4096, 4096 * 3, 4096 * 4096 * 3,
                               128, True, d_hls.data_ptr(), d_l.data_ptr(), lo, hi, d_m.data_ptr(),
                               torch.cuda.current_stream().cuda_stream)
    got = d_hls.cpu().numpy()
    assert np.array_equal(got, want)
    assert np.array_equal(d_l.cpu().numpy(), want[..., 1])
    wm = np.all((want >= np.array(lo, np.uint8)) & (want <= np.array(hi, np.uint8)), axis=-1)
    assert np.array_equal(d_m.cpu().numpy(), wm.astype(np.uint8) * 255)


@pytest.mark.parametrize('shape', [(3, 37, 250), (2, 16, 220), (1, 5, 300), (2, 8, 640), (1, 3, 7)])
def test_convert_hls_widths_and_tail_columns(shape):
    torch = _torch()
    from meterelf_b200 import _engine
    rng = np.random.default_rng(shape[2])
    (n, h, w) = shape
    frames = rng.integers(0, 256, size=(n, h, w, 3), dtype=np.uint8)
    d_in = _dev(frames)
    for s_tail in (True, False):
        d_hls = torch.empty((n, h, w, 3), dtype=torch.uint8, device='cuda')
        _engine.convert_hls_device(0, d_in.data_ptr(), n, h, w, w * 3, h * w * 3, 7, s_tail,
                                   d_hls.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
        got = d_hls.cpu().numpy()
        for f in range(n):
            assert np.array_equal(got[f], oracle.convert_to_hls(frames[f], 7, s_tail))


@pytest.mark.parametrize('set_no', [1, 2])
def test_match_template_map_bit_exact(set_no):
    """TM_CCOEFF map + arg-max on sample crops (BGR input, lightness computed on the fly)."""
    torch = _torch()
    gs = golden_set(set_no)
    crops = gs.subset()['crops'][:6]
    (n, h, w) = crops.shape[:3]
    (rh, rw) = (h - gs.tmpl_h + 1, w - gs.tmpl_w + 1)
    eng = gs.engine(cropped_input=True)
    d_in = _dev(crops)
    d_xy = torch.zeros((n, 2), dtype=torch.int32, device='cuda')
    d_val = torch.zeros((n,), dtype=torch.float32, device='cuda')
    d_map = torch.zeros((n, rh, rw), dtype=torch.float32, device='cuda')
    eng.match_template_device(d_in.data_ptr(), 3, n, h, w, w * 3, h * w * 3, False,
                              d_xy.data_ptr(), d_val.data_ptr(), d_map.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
    got_map = d_map.cpu().numpy()
    for f in range(n):
        light = oracle.convert_to_hls(crops[f], gs.hue_shift)[..., 1]
        (x, y, val, ref_map) = oracle.match_template(np.ascontiguousarray(light), gs.template)
        assert np.array_equal(got_map[f].view(np.uint32), ref_map.view(np.uint32))
        assert tuple(d_xy[f].cpu().numpy()) == (x, y)
        assert np.float32(d_val[f].item()).tobytes() == np.float32(val).tobytes()


@pytest.mark.parametrize('h,w', [(119, 188), (120, 189), (130, 260), (200, 330), (480, 640)])
def test_match_template_gray_sizes_and_ties(h, w):
    """1-channel input, window-size sweep (configs[4]), constant image -> (0, 0)."""
    torch = _torch()
    gs = golden_set(1)
    rng = np.random.default_rng(h * 1000 + w)
    imgs = rng.integers(0, 256, size=(3, h, w), dtype=np.uint8)
    imgs[1] = 77                                      # every score 0 -> first in raster order
    ty, tx = rng.integers(0, h - gs.tmpl_h + 1), rng.integers(0, w - gs.tmpl_w + 1)
    imgs[2, ty:ty + gs.tmpl_h, tx:tx + gs.tmpl_w] = gs.template   # planted template
    eng = gs.engine(cropped_input=True)
    d_in = _dev(imgs)
    (rh, rw) = (h - gs.tmpl_h + 1, w - gs.tmpl_w + 1)
    d_xy = torch.zeros((3, 2), dtype=torch.int32, device='cuda')
    d_val = torch.zeros((3,), dtype=torch.float32, device='cuda')
    d_map = torch.zeros((3, rh, rw), dtype=torch.float32, device='cuda')
    eng.match_template_device(d_in.data_ptr(), 1, 3, h, w, w, h * w, False, d_xy.data_ptr(),
                              d_val.data_ptr(), d_map.data_ptr(), torch.cuda.current_stream().cuda_stream)
    got_map = d_map.cpu().numpy()
    for f in range(3):
        (x, y, val, ref_map) = oracle.match_template(imgs[f], gs.template)
        assert np.array_equal(got_map[f].view(np.uint32), ref_map.view(np.uint32)), f
        assert tuple(d_xy[f].cpu().numpy()) == (x, y), f
    assert tuple(d_xy[1].cpu().numpy()) == (0, 0) and d_val[1].item() == 0.0
    assert tuple(d_xy[2].cpu().numpy()) == (tx, ty)


@pytest.mark.parametrize('tw', [33, 97, 129, 161, 162, 193])
def test_match_template_narrow_templates_on_the_tensor_core_kernel(tw):
    """Templates narrower than 162 px leave whole K-steps of the Toeplitz operand zero; the locate kernel
    neither loads nor multiplies those (``MelfMatchTcArgs::kpp``: 3..7 of 8 K-steps per row pair here, all 8
    from 162 px on).  Result map large enough that the size dispatch takes the Toeplitz kernel."""
    from meterelf_b200 import _utils
    torch = _torch()
    rng = np.random.default_rng(tw)
    (ih, iw, th) = (260, 336 if tw % 2 == 0 else 330, 57)       # 336: rows are whole 16-byte chunks (tensor-copy staging)
    img = rng.integers(0, 256, (ih, iw), dtype=np.uint8)
    tmpl = rng.integers(0, 256, (th, tw), dtype=np.uint8)
    (ty, tx) = (int(rng.integers(0, ih - th + 1)), int(rng.integers(0, iw - tw + 1)))
    planted = img.copy()
    planted[ty:ty + th, tx:tx + tw] = tmpl
    for (image, where) in ((img, None), (planted, (tx, ty))):
        got = _utils.match_template(image, tmpl)
        eng = _utils._template_engine(tmpl, torch.cuda.current_device())
        assert eng.last_locate() == 'melf_match_fused_ov_kernel'
        (x, y, val, _map) = oracle.match_template(image, tmpl, want_map=False)
        assert got.rect == ((x, y), (x + tw, y + th)) and np.float32(got.max_val) == val, tw
        if where:
            assert (x, y) == where


def test_operators_with_the_reference_signatures(tmp_path):
    """``match_template(img, template)`` with ANY template (``_utils.py:91-97``),
    ``get_mask_by_color(hls_image, color, color_range)`` on an HLS image (``_utils.py:113-119``),
    ``ImageFile._find_dials(img_hls)`` and ``ImageFile.get_bgr_image_t()`` (``_image.py:34-66``)."""
    from helpers import write_params_file
    from meterelf_b200 import _colors, _image, _params, _utils
    from meterelf_b200.exceptions import DialsNotFoundError
    rng = np.random.default_rng(12)
    for (ih, iw, th, tw) in ((40, 60, 9, 13), (64, 64, 64, 64), (90, 300, 31, 200), (140, 130, 129, 40)):
        img = rng.integers(0, 256, (ih, iw), dtype=np.uint8)
        tmpl = rng.integers(0, 256, (th, tw), dtype=np.uint8)
        got = _utils.match_template(img, tmpl)
        (x, y, val, _map) = oracle.match_template(img, tmpl, want_map=False)
        assert got.rect == ((x, y), (x + tw, y + th)) and np.float32(got.max_val) == val, (ih, iw, th, tw)
    with pytest.raises(ValueError):
        _utils.match_template(np.zeros((5, 5), np.uint8), np.zeros((6, 3), np.uint8))
    for (h, w) in ((1, 1), (7, 16), (33, 250), (119, 188)):
        hls = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        for (color, crange) in (((125, 80, 130), (9, 45, 35)), ((3, 250, 128), (12, 35, 255)), ((0, 0, 0), (0, 0, 0))):
            (c, r) = (_colors.HlsColor(*color), _colors.HlsColor(*crange))
            got = _utils.get_mask_by_color(hls, c, r)
            (lo, hi) = c.get_range(r)
            want = np.all((hls >= np.array(lo)) & (hls <= np.array(hi)), axis=2).astype(np.uint8) * 255
            assert got.dtype == np.uint8 and np.array_equal(got, want), (h, w, color)
    gs = golden_set(1)
    params = _params.load(write_params_file(gs, tmp_path))
    sub = gs.subset()
    names = [str(n) for n in sub['names']]
    crop = sub['crops'][names.index('20180814215230-01-e136.jpg')]
    imgf = _image.ImageFile('x.jpg', params, crop)
    hls = oracle.convert_to_hls(crop, gs.hue_shift)
    (x, y, val, _map) = oracle.match_template(np.ascontiguousarray(hls[:, :, 1]), gs.template, want_map=False)
    found = imgf._find_dials(hls)
    assert found.rect.top_left == (x, y) and np.float32(found.max_val) == val
    moved = imgf.get_bgr_image_t()
    (dx, dy) = (30 - x, 116 - y)
    want = np.zeros_like(crop)
    (h, w) = crop.shape[:2]
    want[max(0, dy):min(h, h + dy), max(0, dx):min(w, w + dx)] = crop[max(0, -dy):min(h, h - dy), max(0, -dx):min(w, w - dx)]
    assert np.array_equal(moved, want)
    import cv2
    m = np.array([[1, 0, dx], [0, 1, dy]], dtype=np.float32)
    assert np.array_equal(moved, cv2.warpAffine(crop, m, (w, h)))             # what the reference computes
    dark = _image.ImageFile('dark.jpg', params, np.zeros_like(crop))
    with pytest.raises(DialsNotFoundError):
        dark._find_dials(np.zeros_like(crop))


# ---------------------------------------------------------------------------------------
# the Python drop-in layer on top of the C ABI
# ---------------------------------------------------------------------------------------
def test_python_api_image_file_and_errors(tmp_path):
    import cv2
    import yaml
    from meterelf_b200 import _params, _reading
    from meterelf_b200._image import ImageFile
    from meterelf_b200.exceptions import DialsNotFoundError
    gs = golden_set(1)
    cv2.imwrite(str(tmp_path / 'dials.png'), gs.template)
    needle_data = [dict(
        name=gs.names[i], color_range=dict(zip('hls', gs.dials[i]['color_range'])),
        dist_from_center=int(gs.dist_from_center[i]), circle_thickness=int(gs.circle_thickness[i]),
        angle_of_zero=gs.dials[i]['angle_of_zero'], center=list(gs.dials[i]['center']),
        diameter=int(gs.diameters[i]), negative_momentum=gs.dials[i]['negative_momentum'])
        for i in range(4)]
    (x0, y0, x1, y1) = gs.meter_rect
    doc = dict(image_glob='*.png', meter_rect=dict(top_left=[x0, y0], bottom_right=[x1, y1]),
               dials_template='dials.png', dials_template_match_threshold=int(gs.match_threshold),
               dials_template_size=[gs.tmpl_w, gs.tmpl_h], hue_shift=gs.hue_shift,
               needle_color=dict(h=125, l=80, s=130), needle_color_range=dict(h=9, l=45, s=35),
               needle_data=needle_data)
    (tmp_path / 'params.yml').write_text(yaml.safe_dump(doc))
    params = _params.load(str(tmp_path / 'params.yml'))
    sub = gs.subset()
    names = [str(s) for s in sub['names']]
    fn = '20180814215230-01-e136.jpg'
    crop = sub['crops'][names.index(fn)]
    values = _reading.get_meter_value(ImageFile(fn, params, crop))
    assert list(values) == gs.names + ['value']
    assert all(type(v) is float for v in values.values())
    assert '{:07.3f}'.format(values['value']) == '253.623'       # tests/test_meterelf.py:170-188
    assert abs(values['value'] - 253.62306) < 0.000005
    bad = sub['crops'][names.index('20180814021309-01-e01.jpg')]
    with pytest.raises(DialsNotFoundError) as exc:
        _reading.get_meter_value(ImageFile('x.jpg', params, bad))
    assert exc.value.get_message() == 'Dials not found (match val = 0.0)'

    # the CLI on real files (PNG is lossless, so the decoded pixels equal the crops)
    frame = np.zeros((640, 480, 3), np.uint8)
    frame[y0:y1, x0:x1] = crop
    cv2.imwrite(str(tmp_path / 'a.png'), frame)
    from meterelf_b200 import get_meter_values
    out = list(get_meter_values(str(tmp_path / 'params.yml'),
                                [str(tmp_path / 'a.png'), str(tmp_path / 'missing.png')]))
    assert '{:07.3f}'.format(out[0].value) == '253.623' and out[0].error is None
    assert out[1].value is None and out[1].error.get_message() == 'Unable to load image'


def write_params_yml(gs, tmp_path, image_glob='*.jpg'):
    """params.yml + dials template for a golden set, as the reference's sample directories hold."""
    import cv2
    import yaml
    cv2.imwrite(str(tmp_path / 'dials.png'), gs.template)
    nd = len(gs.names)
    needle_data = [dict(
        name=gs.names[i], color_range=dict(zip('hls', gs.dials[i]['color_range'])),
        dist_from_center=int(gs.dist_from_center[i]), circle_thickness=int(gs.circle_thickness[i]),
        angle_of_zero=gs.dials[i]['angle_of_zero'], center=list(gs.dials[i]['center']),
        diameter=int(gs.diameters[i]), negative_momentum=gs.dials[i]['negative_momentum'])
        for i in range(nd)]
    (x0, y0, x1, y1) = gs.meter_rect
    doc = dict(image_glob=image_glob, meter_rect=dict(top_left=[x0, y0], bottom_right=[x1, y1]),
               dials_template='dials.png', dials_template_match_threshold=int(gs.match_threshold),
               dials_template_size=[gs.tmpl_w, gs.tmpl_h], hue_shift=gs.hue_shift,
               needle_color=dict(h=125, l=80, s=130), needle_color_range=dict(h=9, l=45, s=35),
               needle_data=needle_data)
    (tmp_path / 'params.yml').write_text(yaml.safe_dump(doc))
    return str(tmp_path / 'params.yml')


@pytest.mark.parametrize('set_no', [1, 2])
def test_cli_on_jpeg_files_prints_the_golden_lines(set_no, tmp_path, capsys, monkeypatch):
    """``python -m meterelf PARAMS FILES`` (_main.py:8-22) on the sample JPEGs themselves: the files
    are decoded by the ingest kernels (default) and, in a second pass, by cv2.imread on the host
    (MELF_INGEST=host); both must print the reference's golden stdout lines."""
    from test_jpeg_oracle import load_streams
    from meterelf_b200 import _engine, _main
    path = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    streams = load_streams(path if os.path.exists(path) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
    streams = [s for s in streams if s[4] == set_no]
    gs = golden_set(set_no)
    params_file = write_params_yml(gs, tmp_path)
    files = []
    for s in streams:
        files.append(str(tmp_path / s[0]))
        with open(files[-1], 'wb') as fp:
            fp.write(s[1])
    files.append(str(tmp_path / 'missing.jpg'))

    def run():
        _main.main(['meterelf', params_file] + files)
        return capsys.readouterr().out.splitlines()

    def check(lines):
        assert len(lines) == len(files)
        n_exact = 0
        for (line, fn, s) in zip(lines, files, streams):
            exp = gs.expected[s[0]]
            assert line.startswith(fn + ': ')
            text = line[len(fn) + 2:]
            if exp['status'] == 0:
                assert text == exp['line'], s[0]
                n_exact += 1
            else:       # cv2's float32 DFT rounds the match value differently
                assert text.split(' (match val')[0] == exp['line'].split(' (match val')[0], s[0]
        assert lines[-1] == files[-1] + ': UNKNOWN Unable to load image'
        return n_exact

    before = _engine.launch_count()
    gpu_lines = run()
    assert check(gpu_lines) >= 2
    assert _engine.launch_count() > before
    monkeypatch.setenv('MELF_INGEST', 'host')
    try:
        import cv2  # noqa: F401
    except ImportError:
        return
    assert run() == gpu_lines


def test_api_sends_unsupported_files_through_cv2(tmp_path):
    """Files the ingest kernels do not take (progressive JPEG, EXIF-rotated JPEG, a JPEG cut short, PNG)
    go through cv2.imread like in the reference; supported ones in the same call still go to the GPU as
    bytes.  Expected readings: the oracle on what cv2.imread returns for each file."""
    import struct
    import cv2
    from meterelf_b200 import _image, get_meter_values
    gs = golden_set(1)
    params_file = write_params_yml(gs, tmp_path)
    sub = gs.subset()
    names = [str(s) for s in sub['names']]
    crop = sub['crops'][names.index('20180814215230-01-e136.jpg')]
    (x0, y0, x1, y1) = gs.meter_rect
    frame = np.full((640, 480, 3), 90, np.uint8)
    frame[y0:y1, x0:x1] = crop
    files = {k: str(tmp_path / k) for k in ('base.jpg', 'prog.jpg', 'exif.jpg', 'cut.jpg', 'loss.png')}
    cv2.imwrite(files['base.jpg'], frame, [cv2.IMWRITE_JPEG_QUALITY, 95])
    cv2.imwrite(files['prog.jpg'], frame, [cv2.IMWRITE_JPEG_QUALITY, 95, cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    cv2.imwrite(files['loss.png'], frame)
    base = open(files['base.jpg'], 'rb').read()
    tiff = b'II' + struct.pack('<HI', 42, 8) + struct.pack('<H', 1) + \
        struct.pack('<HHIHH', 0x0112, 3, 1, 3, 0) + struct.pack('<I', 0)          # orientation 3: upside down
    seg = b'Exif\0\0' + tiff
    open(files['exif.jpg'], 'wb').write(base[:2] + b'\xff\xe1' + struct.pack('>H', len(seg) + 2) + seg + base[2:])
    open(files['cut.jpg'], 'wb').write(base[:len(base) * 9 // 10])
    routes = {k: _image.gpu_ingest_geometry(open(v, 'rb').read()) is not None for (k, v) in files.items()}
    assert routes == {'base.jpg': True, 'prog.jpg': False, 'exif.jpg': False, 'cut.jpg': False, 'loss.png': False}
    order = list(files)
    got = list(get_meter_values(params_file, [files[k] for k in order]))
    for (k, data) in zip(order, got):
        img = cv2.imread(files[k])
        want = oracle.read_batch(gs.oracle_params(), np.ascontiguousarray(img[y0:y1, x0:x1])[None], n_threads=1)[0]
        line = format_result_line(want, gs.names)
        if int(want['status']) == 0:
            assert data.error is None and data.value == float(want['value']), k
        else:
            assert data.value is None and ('UNKNOWN ' + data.error.get_message()).split(' (match val')[0] == \
                line.split(' (match val')[0], k
    assert got[0].value is not None and got[1].value is not None and got[4].value is not None


def test_debug_overlays_are_written_as_png_files(tmp_path, monkeypatch):
    """DEBUG + MELF_DEBUG_DIR leaves the reference's debug pictures (_reading.py:43-78,
    112-114) as PNG files; what they show is checked against the taps of the GPU path."""
    import cv2
    from meterelf_b200 import _debug, _overlay, _params, _reading, get_meter_values
    gs = golden_set(1)
    params_file = write_params_yml(gs, tmp_path, image_glob='*.png')
    sub = gs.subset()
    names = [str(s) for s in sub['names']]
    crop = sub['crops'][names.index('20180814215230-01-e136.jpg')]
    (x0, y0, x1, y1) = gs.meter_rect
    frame = np.zeros((640, 480, 3), np.uint8)
    frame[y0:y1, x0:x1] = crop
    cv2.imwrite(str(tmp_path / 'a.png'), frame)
    out_dir = tmp_path / 'debug'
    monkeypatch.setattr(_debug, 'DEBUG', {'1'})
    monkeypatch.setenv('MELF_DEBUG_DIR', str(out_dir))
    (data,) = list(get_meter_values(params_file, [str(tmp_path / 'a.png')]))
    assert '{:07.3f}'.format(data.value) == '253.623'
    files = sorted(os.listdir(out_dir))
    assert files == sorted(['a.png.png'] + ['a.png.dial-%s.png' % n for n in gs.names])
    final = cv2.imread(str(out_dir / 'a.png.png'))
    assert final.shape == (2 * gs.tmpl_h, 2 * gs.tmpl_w, 3)
    last = cv2.imread(str(out_dir / ('a.png.dial-%s.png' % gs.names[-1])))
    assert last.shape == (4 * gs.tmpl_h, 4 * gs.tmpl_w, 3)
    # the files hold exactly what the reference showed for this frame (overlay_expected.json)
    with open(os.path.join(GOLDEN, 'overlay_expected.json')) as fp:
        exp = {e['case']: e for e in json.load(fp)}['set1:20180814215230-01-e136.jpg']
    shown = [cv2.imread(str(out_dir / ('a.png.dial-%s.png' % n))) for n in gs.names] + [final]
    assert [hashlib.sha256(np.ascontiguousarray(img).tobytes()).hexdigest() for img in shown] == [
        w['sha256'] for w in exp['windows']]


def test_debug_pictures_equal_what_the_reference_shows(tmp_path):
    """SURVEY 8f-4 pinned to the reference: the arrays the unmodified reference hands to cv2.imshow in
    debug mode (captured by tests/golden/make_golden_overlay.py with cv2.imshow mocked, as the
    reference's own tests do) against _overlay.paint_frame -- per-dial 4x pictures and the final 2x
    picture, sample frames of both sets and synthetic frames ending in every status."""
    from helpers import OVERLAY_CASES, overlay_case_frame, write_params_file
    from meterelf_b200 import _overlay, _params, _reading
    with open(os.path.join(GOLDEN, 'overlay_expected.json')) as fp:
        expected = {e['case']: e for e in json.load(fp)}
    loaded = {}
    n_windows = 0
    seen = set()
    for case in OVERLAY_CASES:
        exp = expected[case['key']]
        if exp['status'] < 0:
            continue                     # the reference's debug mode crashes on this frame (zero momentum)
        key = (case['set'], case['threshold'])
        if key not in loaded:
            d = tmp_path / ('p%d' % len(loaded))
            d.mkdir()
            params = _params.load(write_params_file(golden_set(case['set']), d))
            if case['threshold'] is not None:
                params.dials_match_threshold = case['threshold']
            loaded[key] = (params, _reading.get_engine(params, cropped_input=True))
        (params, eng) = loaded[key]
        frame = overlay_case_frame(case)
        (records, taps) = _overlay.read_with_taps(eng, frame[None])
        assert int(records[0]['status']) == exp['status'], case['key']
        (debug, per_dial) = _overlay.paint_frame(params, eng, frame, records[0], taps['needle_mask'][0],
                                                 taps['counts'][0], taps['momentum'][0])
        shown = [img for (_name, img) in per_dial]
        if exp['status'] == 0:
            shown.append(_overlay.scale_image(debug, 2))
        assert len(shown) == len(exp['windows']), case['key']
        for (k, (img, w)) in enumerate(zip(shown, exp['windows'])):
            assert list(img.shape) == w['shape'], (case['key'], k)
            assert hashlib.sha256(np.ascontiguousarray(img).tobytes()).hexdigest() == w['sha256'], (case['key'], k)
            n_windows += 1
        seen.add(exp['status'])
    assert seen == {0, 2, 3} and n_windows >= 60


@pytest.mark.parametrize('hue_shift,lo,hi', [
    (0, (10, 20, 30), (200, 220, 250)),        # plain interval on every channel
    (128, (116, 35, 95), (134, 125, 165)),     # H interval wraps after removing the shift
    (77, (0, 0, 0), (255, 255, 255)),          # everything
    (200, (30, 0, 0), (29, 255, 255)),         # lo > hi on H: cv2.inRange selects nothing
    (5, (100, 50, 60), (100, 50, 255)),        # single values
    (250, (3, 0, 17), (9, 255, 17)),
])
def test_convert_mask_ranges_bit_exact(hue_shift, lo, hi):
    """Fused inRange of the vector kernel (sign-bit formulation) against numpy on the oracle's HLS."""
    torch = _torch()
    from meterelf_b200 import _engine
    rng = np.random.default_rng(hue_shift)
    frames = rng.integers(0, 256, size=(2, 24, 64, 3), dtype=np.uint8)
    frames[0, :8] = rng.integers(0, 256, size=(8, 64, 1), dtype=np.uint8)        # greys: diff == 0
    d_in = _dev(frames)
    d_m = torch.full((2, 24, 64), 7, dtype=torch.uint8, device='cuda')
    d_l = torch.empty((2, 24, 64), dtype=torch.uint8, device='cuda')
    _engine.convert_hls_device(0, d_in.data_ptr(), 2, 24, 64, 64 * 3, 24 * 64 * 3, hue_shift, True,
                               0, d_l.data_ptr(), lo, hi, d_m.data_ptr(),
                               torch.cuda.current_stream().cuda_stream)
    for f in range(2):
        want = oracle.convert_to_hls(frames[f], hue_shift)
        wm = np.all((want >= np.array(lo, np.uint8)) & (want <= np.array(hi, np.uint8)), axis=-1)
        assert np.array_equal(d_m[f].cpu().numpy(), wm.astype(np.uint8) * 255)
        assert np.array_equal(d_l[f].cpu().numpy(), want[..., 1])


# ---------------------------------------------------------------------------------------
# many passes per persistent CTA, odd item counts, repeated launches: the hand-offs of the
# fused locate kernel (mbarrier phases, per-CTA scratch reuse) and of the dial kernel (per-frame
# slots that reset themselves) must give the same records every time
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize('n_frames', [1, 297, 1500])
def test_large_batches_are_deterministic_and_match_the_oracle(n_frames):
    torch = _torch()
    gs = golden_set(1)
    crops = gs.subset()['crops']
    rng = np.random.default_rng(n_frames)
    idx = rng.integers(0, len(crops), size=n_frames)
    frames = crops[idx].copy()
    # perturb every frame a little so that no two are equal (keeps most of them readable)
    noise = rng.integers(-3, 4, size=frames.shape, dtype=np.int16)
    frames = np.clip(frames.astype(np.int16) + noise, 0, 255).astype(np.uint8)
    from meterelf_b200 import _engine
    eng = gs.engine(cropped_input=True)
    d_in = _dev(frames)
    (h, w) = frames.shape[1:3]
    outs = []
    for _ in range(4):
        d_out = torch.zeros((n_frames, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device='cuda')
        eng.read_device(d_in.data_ptr(), n_frames, h, w, w * 3, h * w * 3, d_out.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        outs.append(d_out.cpu().numpy().tobytes())
    assert all(o == outs[0] for o in outs[1:])
    got = np.frombuffer(outs[0], dtype=_engine.RESULT_DTYPE)
    k = min(n_frames, 96)
    want = oracle.read_batch(gs.oracle_params(), frames[:k], n_threads=8)
    for (g, wnt) in zip(got[:k], want):
        assert_results_equal(g, wnt, len(gs.names))
    assert int(np.sum(got['status'] == 0)) > 0.5 * n_frames


@pytest.mark.parametrize('env_switch', [{'MELF_MATCH_TWO_KERNEL': '1'}, {'MELF_MATCH_OVERLAP': '0'}])
def test_alternative_locate_kernels_still_match(tmp_path, env_switch):
    """The A/B switches (read once per process): MELF_MATCH_TWO_KERNEL=1 = the superseded preparation +
    correlation kernel pair, MELF_MATCH_OVERLAP=0 = the fused kernel with the drain between the passes.
    Checked in a child process against the default (fused, overlapped drain)."""
    import subprocess
    import sys
    gs = golden_set(2)
    crops = gs.subset()['crops'][:10]
    want = gs.engine(cropped_input=True).read_frames(crops)
    np.save(str(tmp_path / 'crops.npy'), crops)
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from helpers import golden_set\n"
        "gs = golden_set(2)\n"
        "out = gs.engine(cropped_input=True).read_frames(np.load(%r))\n"
        "np.save(%r, out)\n"
    ) % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
         os.path.dirname(os.path.abspath(__file__)),
         str(tmp_path / 'crops.npy'), str(tmp_path / 'out.npy'))
    env = dict(os.environ, **env_switch)
    subprocess.run([sys.executable, '-c', code], check=True, env=env, timeout=300)
    got = np.load(str(tmp_path / 'out.npy'))
    assert got.tobytes() == want.tobytes()


def test_small_result_map_path_equals_the_toeplitz_kernel(tmp_path):
    """Sample set 2 (33 x 17 result map) and small single-channel search windows through the positions-as-M /
    frames-as-N GEMM (csrc/melf_match_tc.cu, last section; an opt-in path, MELF_MATCH_SMALL=1, run here in a
    child process) against the default Toeplitz kernel: records and arg-max
    bit-identical -- sample crops, synthetic blob frames (ties, flat frames), a batch that is not a multiple
    of 32, whole 640 x 480 frames with the meter_rect cut on the device, grey windows of three sizes."""
    import subprocess
    import sys
    gs = golden_set(2)
    (x0, y0, x1, y1) = gs.meter_rect
    crops = np.concatenate([gs.subset()['crops'][:40], synthetic_batch(1, 70, y1 - y0, x1 - x0)])[:107]
    rng = np.random.default_rng(3)
    frames = rng.integers(0, 256, size=(37, 640, 480, 3), dtype=np.uint8)
    frames[:, y0:y1, x0:x1] = crops[:37]
    np.save(str(tmp_path / 'crops.npy'), crops)
    np.save(str(tmp_path / 'frames.npy'), frames)
    greys = {}
    for (h, w) in ((135, 220), (119, 188), (130, 200)):
        greys['%dx%d' % (w, h)] = rng.integers(0, 256, size=(45, h, w), dtype=np.uint8)
        np.save(str(tmp_path / ('grey%dx%d.npy' % (w, h))), greys['%dx%d' % (w, h)])
    code = (
        "import sys, numpy as np, torch\n"
        "sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from helpers import golden_set\n"
        "gs = golden_set(2)\n"
        "d = %r\n"
        "np.save(d + '/out_crops.npy', gs.engine(cropped_input=True, threshold=-1e30).read_frames(np.load(d + '/crops.npy')))\n"
        "np.save(d + '/out_frames.npy', gs.engine(cropped_input=False).read_frames(np.load(d + '/frames.npy')))\n"
        "eng = gs.engine(cropped_input=True)\n"
        "for name in ('220x135', '188x119', '200x130'):\n"
        "    g = torch.from_numpy(np.load(d + '/grey%%s.npy' %% name)).cuda()\n"
        "    (n, h, w) = g.shape\n"
        "    xy = torch.zeros((n, 2), dtype=torch.int32, device='cuda'); val = torch.zeros((n,), dtype=torch.float32, device='cuda')\n"
        "    eng.match_template_device(g.data_ptr(), 1, n, h, w, w, h * w, False, xy.data_ptr(), val.data_ptr())\n"
        "    torch.cuda.synchronize()\n"
        "    np.save(d + '/out_grey%%s.npy' %% name, np.concatenate([xy.cpu().numpy().astype(np.float64), val.cpu().numpy().astype(np.float64)[:, None]], axis=1))\n"
    ) % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)), str(tmp_path))
    results = {}
    for (tag, env_switch) in (('small', {'MELF_MATCH_SMALL': '1'}), ('toeplitz', {'MELF_MATCH_SMALL': '0'})):
        subprocess.run([sys.executable, '-c', code], check=True, env=dict(os.environ, **env_switch), timeout=600)
        results[tag] = {k: np.load(str(tmp_path / k)) for k in ('out_crops.npy', 'out_frames.npy', 'out_grey220x135.npy',
                                                                'out_grey188x119.npy', 'out_grey200x130.npy')}
    for k in results['small']:
        assert results['small'][k].tobytes() == results['toeplitz'][k].tobytes(), k
    want = oracle.read_batch(gs.oracle_params(threshold=-1e30), crops, n_threads=8)
    for (g, w) in zip(results['small']['out_crops.npy'], want):
        assert_results_equal(g, w, len(gs.names))
    for (name, imgs) in greys.items():
        for i in (0, 7, 44):
            (x, y, val, _m) = oracle.match_template(imgs[i], gs.template, want_map=False)
            assert tuple(results['small']['out_grey%s.npy' % name][i]) == (float(x), float(y), float(val)), (name, i)
