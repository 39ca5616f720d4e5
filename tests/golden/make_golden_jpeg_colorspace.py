#!/usr/bin/env python
"""cv2.imdecode (libjpeg) on streams re-labelled to take every branch of libjpeg's colour space
choice (``tests/helpers.jpeg_colour_space_variants``): 3-component baseline files are YCbCr unless
an Adobe APP14 marker says transform 0 or the component ids are 'R','G','B' (and no JFIF marker
overrides either) -- then libjpeg copies the components as R, G, B.

    python tests/golden/make_golden_jpeg_colorspace.py   ->  tests/golden/jpeg_colorspace.json

The base streams are three committed sample streams of ``jpeg_subset.npz`` (4:2:0) plus one
4:4:4 and one 4:2:2 synthetic stream from the same file."""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from helpers import jpeg_colour_space_bases, jpeg_colour_space_variants  # noqa: E402


def main():
    out = []
    for (name, data) in jpeg_colour_space_bases():
        for (key, variant, is_rgb) in jpeg_colour_space_variants(data):
            img = cv2.imdecode(np.frombuffer(variant, np.uint8), cv2.IMREAD_COLOR)
            assert img is not None, (name, key)
            out.append({'name': name, 'variant': key, 'rgb': is_rgb, 'shape': list(img.shape[:2]),
                        'sha256': hashlib.sha256(img.tobytes()).hexdigest()})
    with open(os.path.join(HERE, 'jpeg_colorspace.json'), 'w') as fp:
        json.dump(out, fp, indent=0)
    print(len(out), 'vectors;', sum(v['rgb'] for v in out), 'read as RGB')


if __name__ == '__main__':
    main()
