#!/usr/bin/env python
"""Golden vectors of the ingest row: JPEG bytes -> what ``cv2.imread`` returns.

    python tests/golden/make_golden_jpeg.py

Run in the build container (reads the reference's sample JPEGs and uses the container's cv2,
i.e. OpenCV's bundled libjpeg-turbo -- the decoder ``meterelf/_image.py:49`` calls).  Writes

* ``tests/golden/jpeg_subset.npz`` (committed): the bytes of a few sample JPEGs of both sets,
  plus small synthetic streams that cv2 itself encoded from crops of one sample (4:4:4, 4:2:2,
  4:2:0, grey, restart intervals, odd and tiny sizes); for every stream the SHA-256 of the BGR
  array ``cv2.imdecode`` returns, and for the sample files the expected reading line.
* ``tests/golden/_full/jpeg_all.npz`` (git-ignored, travels with the gpurun snapshot): the bytes
  of all 304 sample JPEGs with the same digests, in golden-stdout order per set.

The JPEG files are camera data, not reference source.
"""
import glob
import hashlib
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference'
SUBSET_PER_SET = 4


def digest(arr):
    return hashlib.sha256(np.ascontiguousarray(arr).tobytes()).hexdigest()


def pack(streams):
    """[(name, bytes, set_no)] -> arrays for np.savez."""
    blob = np.frombuffer(b''.join(s[1] for s in streams), np.uint8)
    sizes = np.array([len(s[1]) for s in streams], np.int64)
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    shapes = []
    digests = []
    for (_, data, _) in streams:
        img = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
        shapes.append(img.shape[:2])
        digests.append(digest(img))
    return dict(blob=blob, offsets=offsets, names=np.array([s[0] for s in streams]),
                set_no=np.array([s[2] for s in streams], np.int32),
                shapes=np.array(shapes, np.int32), sha256=np.array(digests))


def synthetic(src):
    rng = np.random.default_rng(20181001)
    samp = {'444': cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444, '422': cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422,
            '420': cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420}
    out = []
    for (hh, ww) in [(1, 1), (3, 5), (16, 16), (17, 33), (40, 6), (64, 48), (119, 188)]:
        for (sname, sv) in samp.items():
            for (rst, q, grey) in [(0, 75, False), (3, 100, False), (0, 90, True)]:
                img = src[150:150 + hh, 120:120 + ww]
                if grey:
                    if sname != '444':
                        continue
                    img = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
                if q == 100:      # noisy + near-lossless: exercises clamping and long codes
                    img = (img.astype(int) + rng.integers(-90, 90, img.shape)).clip(0, 255).astype(np.uint8)
                (ok, buf) = cv2.imencode('.jpg', img, [
                    cv2.IMWRITE_JPEG_QUALITY, q, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sv,
                    cv2.IMWRITE_JPEG_RST_INTERVAL, rst])
                assert ok
                out.append(('synthetic-%dx%d-%s-rst%d-q%d%s' % (ww, hh, sname, rst, q, '-grey' if grey else ''),
                            buf.tobytes(), 0))
    return out


def main():
    full = []
    subset = []
    for set_no in (1, 2):
        files = sorted(glob.glob(os.path.join(REF, 'sample-images%d' % set_no, '*.jpg')))
        streams = [(os.path.basename(fn), open(fn, 'rb').read(), set_no) for fn in files]
        for (name, data, _) in streams:          # the digest really is cv2.imread's
            assert digest(cv2.imread(os.path.join(REF, 'sample-images%d' % set_no, name))) == digest(
                cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR))
        full += streams
        stride = max(1, len(streams) // SUBSET_PER_SET)
        subset += streams[::stride][:SUBSET_PER_SET]
    # (not 20180814021309-01-e01.jpg: that frame is flat grey, its q75 streams were all-DC)
    src = cv2.imread(os.path.join(REF, 'sample-images1', '20180814215230-01-e136.jpg'))
    subset += synthetic(src)
    np.savez_compressed(os.path.join(HERE, 'jpeg_subset.npz'), **pack(subset))
    os.makedirs(os.path.join(HERE, '_full'), exist_ok=True)
    np.savez(os.path.join(HERE, '_full', 'jpeg_all.npz'), **pack(full))
    print('subset: %d streams, full: %d streams' % (len(subset), len(full)))


if __name__ == '__main__':
    main()
