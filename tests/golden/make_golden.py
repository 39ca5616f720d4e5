#!/usr/bin/env python
"""Generate the golden fixtures by running the UNMODIFIED reference in this container.

    python tests/golden/make_golden.py [--full-only]

Reads ``/root/reference`` (sample JPEGs, params, golden stdout) through
``tools/ref_harness.py`` (yaml/findContours shims, IPP off) and writes

* ``tests/golden/set{1,2}_expected.json`` (committed): for EVERY sample frame the golden
  stdout line plus the reference's numeric results (status, match location, cv2's
  ``max_val``, float64 dial positions and value as hex floats).
* ``tests/golden/set{1,2}_subset.npz`` (committed): a subset of frames with their
  decoded ``meter_rect`` crops (lossless) and per-stage taps of the reference: dial
  colour, inRange mask, closed mask, needle mask, needle/outer point counts.
* ``tests/golden/set{1,2}_params.npz`` (committed): the dial template and params as
  plain arrays, so no test needs the reference tree.
* ``tests/golden/_full/set{1,2}_crops.npz`` (git-ignored, travels to the GPU box with
  the gpurun snapshot): the crops of ALL frames, for the full 304-line readings diff.

The decoded crops are JPEG-decoder output (libjpeg-turbo via cv2.imread), i.e. data,
not reference source.
"""
import argparse
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))

import ref_harness  # noqa: E402

SUBSET_ALWAYS = {
    1: ['20180814021309-01-e01.jpg', '20180814021310-00-e02.jpg',
        '20180814215230-01-e136.jpg'],
    2: ['20181001_113527-01-e14.jpg'],      # reading 9.9e-9 below a print boundary
}
SUBSET_STRIDE = {1: 4, 2: 7}
HLS_TAP_FRAMES = 2


def run_reference(set_no, want_taps_for):
    import cv2
    ref_harness.load()
    from meterelf import _image, _params, _reading
    from meterelf.exceptions import (
        DialAngleDeterminingError, DialsNotFoundError, ImageProcessingError,
        NeedleContoursNotFoundError)

    sdir = ref_harness.sample_dir(set_no)
    params = _params.load(os.path.join(sdir, 'params.yml'))
    names = list(params.dial_centers.keys())
    golden_lines = dict(
        line.split(': ', 1) for line in ref_harness.golden_stdout(set_no).splitlines())

    captured = {}
    orig_color = _reading.get_dial_color
    orig_mask = _reading.get_mask_by_color
    orig_needle = _reading.get_needle_points
    orig_erode = cv2.erode

    def color_tap(dials_hls, dial_data):
        c = orig_color(dials_hls, dial_data)
        captured.setdefault('color', {})[dial_data.name] = [int(v) for v in c]
        captured['cur'] = dial_data.name
        return c

    def mask_tap(hls, color, color_range):
        m = orig_mask(hls, color, color_range)
        captured.setdefault('mask_orig', {})[captured['cur']] = m.copy()
        return m

    def erode_tap(*a, **k):
        m = orig_erode(*a, **k)
        captured.setdefault('mask_de', {})[captured['cur']] = m.copy()
        return m

    def needle_tap(p, dials_hls, dial_data, debug):
        (pts, mask) = orig_needle(p, dials_hls, dial_data, debug)
        captured.setdefault('needle_mask', {})[dial_data.name] = mask.copy()
        captured.setdefault('n_needle', {})[dial_data.name] = len(pts)
        outer = cv2.findNonZero(mask & dial_data.circle_mask)
        captured.setdefault('n_outer', {})[dial_data.name] = 0 if outer is None else len(outer)
        return (pts, mask)

    _reading.get_dial_color = color_tap
    _reading.get_mask_by_color = mask_tap
    _reading.get_needle_points = needle_tap
    cv2.erode = erode_tap

    (x0, y0) = params.meter_rect.top_left
    (x1, y1) = params.meter_rect.bottom_right
    expected = {}
    crops = {}
    taps = {}
    try:
        for fn in ref_harness.sample_files(set_no):
            img = cv2.imread(os.path.join(sdir, fn))
            crop = np.ascontiguousarray(img[y0:y1, x0:x1])
            crops[fn] = crop
            captured.clear()
            imgf = _image.ImageFile(fn, params, crop)
            hls = imgf.get_hls_image()
            from meterelf._utils import match_template
            m = match_template(cv2.split(hls)[1], _image._get_dials_template(params))
            rec = {'line': golden_lines[fn], 'match_x': int(m.rect.top_left[0]),
                   'match_y': int(m.rect.top_left[1]), 'cv2_max_val': float(m.max_val)}
            try:
                values = _reading.get_meter_value(imgf)
                rec['status'] = 0
                rec['dial_pos'] = [float(values[k]).hex() for k in names]
                rec['value'] = float(values['value']).hex()
                line = '{:07.3f}'.format(values['value'])
            except ImageProcessingError as e:
                rec['status'] = {DialsNotFoundError: 1, NeedleContoursNotFoundError: 2,
                                 DialAngleDeterminingError: 3}[type(e)]
                line = 'UNKNOWN ' + e.get_message()
            assert line == golden_lines[fn], (fn, line, golden_lines[fn])
            expected[fn] = rec
            if fn in want_taps_for and rec['status'] == 0:
                taps[fn] = {k: dict(v) for (k, v) in captured.items() if k != 'cur'}
                taps[fn]['hls'] = np.asarray(hls).copy()
    finally:
        _reading.get_dial_color = orig_color
        _reading.get_mask_by_color = orig_mask
        _reading.get_needle_points = orig_needle
        cv2.erode = orig_erode
    return (params, names, expected, crops, taps)


def save_params(set_no, params, names):
    import cv2
    from meterelf._dial_data import get_dial_data
    tmpl = cv2.imread(params.dials_file, cv2.IMREAD_GRAYSCALE)
    dd = get_dial_data(params)
    np.savez_compressed(
        os.path.join(HERE, 'set%d_params.npz' % set_no),
        template=tmpl,
        names=np.array(names),
        meter_rect=np.array(params.meter_rect.top_left + params.meter_rect.bottom_right),
        hue_shift=params.hue_shift,
        match_threshold=params.dials_match_threshold,
        centers=np.array([dd[k].center for k in names], np.float64),
        diameters=np.array([params.dial_centers[k].diameter for k in names]),
        dist_from_center=np.array([params.needle_dists_from_dial_center[k] for k in names]),
        circle_thickness=np.array([params.needle_circle_mask_thickness[k] for k in names]),
        color_range=np.array([[int(c) for c in params.dial_color_range[k]] for k in names]),
        angle_of_zero=np.array([params.needle_angles_of_zero[k] for k in names], np.float64),
        negative_momentum=np.array([k in params.negative_momentum_dials for k in names]),
        ref_mask=np.packbits(np.stack([dd[k].mask for k in names]) > 0, axis=-1),
        ref_circle_mask=np.packbits(np.stack([dd[k].circle_mask for k in names]) > 0, axis=-1),
    )


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--full-only', action='store_true',
                    help='only (re)write the git-ignored full crop sets')
    args = ap.parse_args()
    os.makedirs(os.path.join(HERE, '_full'), exist_ok=True)
    for set_no in (1, 2):
        files = ref_harness.sample_files(set_no)
        subset = sorted(set(files[::SUBSET_STRIDE[set_no]]) | set(SUBSET_ALWAYS[set_no]))
        (params, names, expected, crops, taps) = run_reference(set_no, set(subset))
        np.savez_compressed(
            os.path.join(HERE, '_full', 'set%d_crops.npz' % set_no),
            names=np.array(files), crops=np.stack([crops[f] for f in files]))
        if args.full_only:
            continue
        with open(os.path.join(HERE, 'set%d_expected.json' % set_no), 'w') as fp:
            json.dump({'dial_names': names, 'frames': expected}, fp, indent=0, sort_keys=True)
        save_params(set_no, params, names)
        tapped = [f for f in subset if f in taps]

        def stack_masks(key):
            return np.packbits(np.stack([
                np.stack([taps[f][key][k] for k in names]) for f in tapped]) > 0, axis=-1)
        np.savez_compressed(
            os.path.join(HERE, 'set%d_subset.npz' % set_no),
            names=np.array(subset),
            crops=np.stack([crops[f] for f in subset]),
            tap_names=np.array(tapped),
            tap_color=np.array([[taps[f]['color'][k] for k in names] for f in tapped]),
            tap_mask_orig=stack_masks('mask_orig'),
            tap_mask_de=stack_masks('mask_de'),
            tap_needle_mask=stack_masks('needle_mask'),
            tap_n_needle=np.array([[taps[f]['n_needle'][k] for k in names] for f in tapped]),
            tap_n_outer=np.array([[taps[f]['n_outer'][k] for k in names] for f in tapped]),
            tap_hls_names=np.array(tapped[:HLS_TAP_FRAMES]),
            tap_hls=np.stack([taps[f]['hls'] for f in tapped[:HLS_TAP_FRAMES]]),
        )
        print('set', set_no, 'frames', len(files), 'subset', len(subset), 'tapped', len(tapped))


if __name__ == '__main__':
    main()
