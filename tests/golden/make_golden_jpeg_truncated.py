#!/usr/bin/env python
"""Golden vectors for JPEG files that are cut short (run in the build container, needs cv2).

    python tests/golden/make_golden_jpeg_truncated.py

For every committed stream WITHOUT restart intervals (tests/golden/jpeg_subset.npz) the first
``keep`` bytes are written to a file and read back with ``cv2.imread`` -- the reference's own
call (meterelf/_image.py:49).  libjpeg finishes the MCU in which the data runs dry on zero bits
and leaves the rest of the image mid-grey (jdhuff.c insufficient_data); the SHA-256 of what it
returns goes to ``tests/golden/jpeg_truncated.json``.  Streams with restart intervals are left
out: there libjpeg additionally resynchronises on the fake EOI its file source appends, which
neither the oracle nor the kernels restate.
"""
import hashlib
import json
import os
import sys
import tempfile

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from test_jpeg_oracle import load_streams  # noqa: E402

FRACTIONS = (0.12, 0.35, 0.5, 0.77, 0.93, 0.999)


def main():
    streams = load_streams(os.path.join(HERE, 'jpeg_subset.npz'))
    out = []
    with tempfile.TemporaryDirectory() as tmp:
        for (name, data, shape, _sha, _set) in streams:
            if '-rst' in name and '-rst0' not in name:      # restart interval > 0
                continue
            for frac in FRACTIONS:
                keep = int(len(data) * frac)
                path = os.path.join(tmp, 'cut.jpg')
                with open(path, 'wb') as fp:
                    fp.write(data[:keep])
                img = cv2.imread(path)
                if img is None:                       # cut inside the headers: nothing to compare
                    continue
                assert img.shape == tuple(shape) + (3,)
                out.append({'name': name, 'keep': keep,
                            'sha256': hashlib.sha256(np.ascontiguousarray(img).tobytes()).hexdigest()})
    with open(os.path.join(HERE, 'jpeg_truncated.json'), 'w') as fp:
        json.dump(out, fp, indent=0)
    print('%d truncated cases from %d streams' % (len(out), len({c['name'] for c in out})))


if __name__ == '__main__':
    main()
