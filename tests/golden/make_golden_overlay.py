#!/usr/bin/env python
"""What the UNMODIFIED reference hands to ``cv2.imshow`` in debug mode (``_reading.py:43-78, 112-114,
143-144``), captured by mocking ``cv2.imshow`` like the reference's own tests do
(``tests/test_meterelf.py:19-28``); no display is involved.

    python tests/golden/make_golden_overlay.py   ->  tests/golden/overlay_expected.json

Frames: sample crops of both sets (committed subset) and synthetic blob frames that end in every
status (``tests/helpers.OVERLAY_CASES``).  Per frame: status and, per window in call order, the title
suffix, the array shape and the SHA-256 of its bytes.  The generator also checks the product's
``hls_to_bgr_full`` / ``resize_linear`` against cv2 on these very frames.
"""
import hashlib
import json
import os
import sys
import tempfile
from unittest.mock import patch

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import ref_harness  # noqa: E402
from helpers import OVERLAY_CASES, overlay_case_frame  # noqa: E402
from make_golden_synthetic import params_file  # noqa: E402


def main():
    import cv2
    ref_harness.load()
    from meterelf import _debug, _image, _params, _reading
    from meterelf._utils import convert_to_bgr
    from meterelf.exceptions import ImageProcessingError
    from meterelf_b200 import _overlay
    out = []
    with tempfile.TemporaryDirectory() as tmp:
        loaded = {}
        for case in OVERLAY_CASES:
            set_no = case['set']
            if set_no not in loaded:
                loaded[set_no] = _params.load(params_file(set_no, tmp))
            params = loaded[set_no]
            params.dials_match_threshold = case['threshold'] if case['threshold'] is not None else (
                _params.load(params_file(set_no, tmp)).dials_match_threshold)
            frame = overlay_case_frame(case)
            shown = []

            def imshow(title, img):
                shown.append((title, np.array(img)))
            status = 0
            with patch.object(_debug, 'DEBUG', new={'1'}), patch('cv2.imshow', imshow), patch('cv2.waitKey'):
                imgf = _image.ImageFile('dir/frame.jpg', params, frame)
                try:
                    _reading.get_meter_value(imgf)
                except ImageProcessingError as e:
                    status = {'DialsNotFoundError': 1, 'NeedleContoursNotFoundError': 2,
                              'DialAngleDeterminingError': 3}[type(e).__name__]
                except ValueError:
                    # the reference's debug mode divides by a zero momentum (a dial without needle
                    # points) and dies on NaN -> int (_reading.py:44-49): nothing to compare with
                    status = -1
                if status != 1:
                    dials_hls = _image.ImageFile('x', params, frame).get_dials_hls()
                    back = dials_hls - np.array([params.hue_shift, 0, 0], np.uint8)
                    assert np.array_equal(_overlay.hls_to_bgr_full(back), convert_to_bgr(params, dials_hls)), case
            for (_t, img) in shown:
                small = img[::4, ::4] if img.shape[0] > 300 else img[::2, ::2]
                for s in (2, 4):
                    assert np.array_equal(_overlay.resize_linear(small, s),
                                          cv2.resize(small, (small.shape[1] * s, small.shape[0] * s))), case
            out.append({'case': case['key'], 'status': status, 'windows': [
                {'title': t, 'shape': list(img.shape), 'sha256': hashlib.sha256(img.tobytes()).hexdigest()}
                for (t, img) in shown]})
            print(case['key'], 'status', status, 'windows', len(shown))
    with open(os.path.join(HERE, 'overlay_expected.json'), 'w') as fp:
        json.dump(out, fp, indent=0)


if __name__ == '__main__':
    main()
