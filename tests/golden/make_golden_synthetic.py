#!/usr/bin/env python
"""Golden fixtures for the branches the sample images never reach, made by running the
UNMODIFIED reference (``tools/ref_harness.py``) on synthetic blob frames.

    python tests/golden/make_golden_synthetic.py

The frames are not stored: they come from ``tests/helpers.synthetic_blob_frames`` with the seeds
below (the very batches ``tests/test_gpu_parity.py::test_synthetic_blob_frames_all_branches``
reads), and the fixture keeps a SHA-256 of each batch so that a generator drift is noticed.
The match threshold is forced to -1e30 so that every frame reaches the dial stage.

Cases: sample params of set 1 and set 2, and "set 3" = set 2's params with the dial centres
moved so that every disc is cut by the frame of the dials rect (left, bottom, right, top):
``cv2.findContours`` pads the image, so components on the frame are contours like any other
(``_reading.py:132-135``).  Set 3's params are written to ``set3_params.npz`` like the others.

Per frame the fixture holds what the reference computed: match location and cv2's max_val
(``_image.py:57-66``), status with the failing dial (``_reading.py:137-138``) or the unreadable
dials (``_reading.py:98-106``), the float64 dial positions and value as hex floats, and per dial
the number of needle points, the area of the selected contour (``_reading.py:140-148``) and the
number of ring points of the needle (``_reading.py:43``).

Writes ``tests/golden/synthetic_expected.json`` and ``tests/golden/set3_params.npz``.
"""
import hashlib
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import ref_harness  # noqa: E402
from helpers import SYNTHETIC_CASES as CASES, synthetic_batch  # noqa: E402
from make_golden import save_params  # noqa: E402

SET3_CENTERS = {'0.0001': [12.3, 63.4], '0.001': [94.0, 105.0], '0.01': [174.5, 71.7], '0.1': [160.9, 9.5]}


def params_file(set_no, tmp):
    """params.yml of a set; set 3 is written to `tmp` from set 2's with the centres moved."""
    if set_no in (1, 2):
        return os.path.join(ref_harness.sample_dir(set_no), 'params.yml')
    import yaml
    src = ref_harness.sample_dir(2)
    with open(os.path.join(src, 'params.yml')) as fp:
        doc = yaml.safe_load(fp)
    for needle in doc['needle_data']:
        needle['center'] = SET3_CENTERS[needle['name']]
    shutil.copy(os.path.join(src, doc['dials_template']), os.path.join(tmp, doc['dials_template']))
    path = os.path.join(tmp, 'params.yml')
    with open(path, 'w') as fp:
        yaml.safe_dump(doc, fp)
    return path


def run_case(params, names, frames):
    import cv2
    from meterelf import _image, _reading
    from meterelf._utils import match_template
    from meterelf.exceptions import (
        DialAngleDeterminingError, DialsNotFoundError, ImageProcessingError,
        NeedleContoursNotFoundError)

    cur = {}
    orig_needle = _reading.get_needle_points
    orig_area = cv2.contourArea

    def needle_tap(p, dials_hls, dial_data, debug):
        cur['areas'] = []
        try:
            (pts, mask) = orig_needle(p, dials_hls, dial_data, debug)
        finally:
            cur.setdefault('area', {})[dial_data.name] = max(cur['areas']) if cur['areas'] else -1.0
        cur.setdefault('n_needle', {})[dial_data.name] = len(pts)
        outer = cv2.findNonZero(mask & dial_data.circle_mask)
        cur.setdefault('n_outer', {})[dial_data.name] = 0 if outer is None else len(outer)
        return (pts, mask)

    def area_tap(contour, *a, **k):
        v = orig_area(contour, *a, **k)
        cur['areas'].append(float(v))
        return v

    _reading.get_needle_points = needle_tap
    cv2.contourArea = area_tap
    out = []
    try:
        for (i, frame) in enumerate(frames):
            cur.clear()
            imgf = _image.ImageFile('frame%d' % i, params, frame)
            hls = imgf.get_hls_image()
            m = match_template(cv2.split(hls)[1], _image._get_dials_template(params))
            rec = {'match_x': int(m.rect.top_left[0]), 'match_y': int(m.rect.top_left[1]),
                   'cv2_max_val': float(m.max_val)}
            try:
                values = _reading.get_meter_value(imgf)
                rec['status'] = 0
                rec['dial_pos'] = [float(values[k]).hex() for k in names]
                rec['value'] = float(values['value']).hex()
            except ImageProcessingError as e:
                rec['status'] = {DialsNotFoundError: 1, NeedleContoursNotFoundError: 2,
                                 DialAngleDeterminingError: 3}[type(e)]
                rec['message'] = e.get_message()
                if rec['status'] == 2:
                    rec['failed_dial'] = names.index(e.extra_info['dial'])
                if rec['status'] == 3:
                    bad = e.extra_info['unreadable dials'].split(', ')
                    rec['unreadable_mask'] = sum(1 << names.index(b) for b in bad)
            for key in ('n_needle', 'area', 'n_outer'):
                rec[key] = [cur.get(key, {}).get(k) for k in names]
            out.append(rec)
    finally:
        _reading.get_needle_points = orig_needle
        cv2.contourArea = orig_area
    return out


def main():
    ref_harness.load()
    from meterelf import _params
    doc = {}
    with tempfile.TemporaryDirectory() as tmp:
        loaded = {}
        for (key, set_no, seed, n) in CASES:
            if set_no not in loaded:
                params = _params.load(params_file(set_no, tmp))
                if set_no == 3:
                    save_params(3, params, list(params.dial_centers.keys()))
                params.dials_match_threshold = -1e30
                loaded[set_no] = params
            params = loaded[set_no]
            names = list(params.dial_centers.keys())
            (x0, y0) = params.meter_rect.top_left
            (x1, y1) = params.meter_rect.bottom_right
            frames = synthetic_batch(seed, n, y1 - y0, x1 - x0)
            records = run_case(params, names, frames)
            doc[key] = {'set': set_no, 'seed': seed, 'n': n, 'dial_names': names,
                        'sha256': hashlib.sha256(frames.tobytes()).hexdigest(), 'frames': records}
            hist = {}
            for r in records:
                hist[r['status']] = hist.get(r['status'], 0) + 1
            small = sum(1 for r in records for a in r['area'] if a is not None and 0 <= a <= 100)
            print(key, 'statuses', hist, 'dials on the <=100 branch', small)
    with open(os.path.join(HERE, 'synthetic_expected.json'), 'w') as fp:
        json.dump(doc, fp, indent=0, sort_keys=True)


if __name__ == '__main__':
    main()
