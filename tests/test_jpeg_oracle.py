"""The JPEG oracle (oracle/jpeg_oracle.c) against what cv2.imread returned in the build
container for the reference's sample files and for synthetic streams."""
import hashlib
import os

import numpy as np
import pytest

import oracle
from helpers import GOLDEN


def load_streams(path):
    d = np.load(path)
    blob = d['blob'].tobytes()
    off = d['offsets']
    return [(str(d['names'][i]), blob[off[i]:off[i + 1]], tuple(int(v) for v in d['shapes'][i]),
             str(d['sha256'][i]), int(d['set_no'][i])) for i in range(len(d['names']))]


def test_oracle_matches_cv2_on_committed_streams():
    streams = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    assert len(streams) >= 50
    for (name, data, shape, sha, _) in streams:
        img = oracle.decode_jpeg(data)
        assert img.shape == shape + (3,), name
        assert hashlib.sha256(img.tobytes()).hexdigest() == sha, name


def test_oracle_matches_cv2_on_every_sample_file():
    path = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    if not os.path.exists(path):
        pytest.skip('full JPEG set not present (git-ignored; made by make_golden_jpeg.py)')
    streams = load_streams(path)
    assert len(streams) == 304
    for (name, data, shape, sha, _) in streams[::3]:
        assert hashlib.sha256(oracle.decode_jpeg(data).tobytes()).hexdigest() == sha, name


def test_oracle_rejects_what_it_does_not_restate():
    (name, data, _, _, _) = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))[0]
    with pytest.raises(oracle.JpegError):
        oracle.decode_jpeg(b'not a jpeg')
    with pytest.raises(oracle.JpegError):
        oracle.decode_jpeg(data[:200])                       # headers cut short
    prog = bytearray(data)
    i = prog.find(b'\xff\xc0')
    prog[i + 1] = 0xC2                                       # claim progressive
    with pytest.raises(oracle.JpegError):
        oracle.decode_jpeg(bytes(prog))


def test_oracle_matches_cv2_on_files_cut_short():
    """jdhuff.c insufficient_data: the MCU in which the data runs dry is finished on zero bits, the
    rest stays mid-grey.  Vectors: cv2.imread on truncated files (make_golden_jpeg_truncated.py)."""
    import json
    streams = {s[0]: s[1] for s in load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))}
    with open(os.path.join(GOLDEN, 'jpeg_truncated.json')) as fp:
        cases = json.load(fp)
    assert len(cases) >= 80
    for case in cases:
        img = oracle.decode_jpeg(streams[case['name']][:case['keep']])
        assert hashlib.sha256(img.tobytes()).hexdigest() == case['sha256'], (case['name'], case['keep'])


def test_colour_space_choice_follows_libjpeg():
    """3-component files are RGB, not YCbCr, for libjpeg when an Adobe APP14 marker says
    transform 0 or the component ids are 'R','G','B' (no JFIF marker).  The oracle decodes every
    variant like cv2.imdecode did (make_golden_jpeg_colorspace.py); the product's header probe
    accepts exactly the YCbCr ones, so the RGB ones take the cv2.imread route."""
    import json
    from helpers import jpeg_colour_space_bases, jpeg_colour_space_variants
    from meterelf_b200 import _engine, _image
    with open(os.path.join(GOLDEN, 'jpeg_colorspace.json')) as fp:
        vectors = {(v['name'], v['variant']): v for v in json.load(fp)}
    n_rgb = 0
    for (name, data) in jpeg_colour_space_bases():
        base_sha = None
        for (key, variant, is_rgb) in jpeg_colour_space_variants(data):
            v = vectors[(name, key)]
            assert v['rgb'] == is_rgb
            img = oracle.decode_jpeg(variant)
            assert hashlib.sha256(img.tobytes()).hexdigest() == v['sha256'], (name, key)
            if key == 'jfif':
                base_sha = v['sha256']
            assert (v['sha256'] == base_sha) == (not is_rgb), (name, key)   # the RGB ones really differ
            geometry = _image.gpu_ingest_geometry(variant)
            if is_rgb:
                n_rgb += 1
                assert geometry is None, (name, key)
                with pytest.raises(_engine.EngineError) as exc:
                    _engine.jpeg_info(variant)
                assert exc.value.code == _engine.ERR_UNSUPPORTED and 'RGB' in str(exc.value)
            else:
                assert geometry is not None and geometry[:2] == tuple(v['shape']), (name, key)
    assert n_rgb == 8


def fresh_encodes(seed=1, sizes=((16, 16), (33, 47), (120, 200), (481, 643))):
    """(tag, stream) of synthetic images encoded here by cv2 / libjpeg: three qualities x {standard tables,
    optimised Huffman tables, restart interval 5} x {4:2:0, 4:4:4 where this cv2 can ask for it}."""
    import cv2
    rng = np.random.default_rng(seed)
    out = []
    for (h, w) in sizes:
        base = rng.integers(0, 256, (h // 8 + 2, w // 8 + 2, 3), dtype=np.uint8)
        img = cv2.resize(base, (w, h), interpolation=cv2.INTER_CUBIC)
        img = np.clip(img.astype(int) + rng.integers(-12, 13, img.shape), 0, 255).astype(np.uint8)
        for q in (35, 75, 95):
            for (label, extra) in (('std', []), ('opt', [cv2.IMWRITE_JPEG_OPTIMIZE, 1]),
                                   ('rst', [cv2.IMWRITE_JPEG_RST_INTERVAL, 5])):
                samplings = [('420', [])]
                if hasattr(cv2, 'IMWRITE_JPEG_SAMPLING_FACTOR'):
                    samplings.append(('444', [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444]))
                for (sname, sflags) in samplings:
                    (ok, buf) = cv2.imencode('.jpg', img, [cv2.IMWRITE_JPEG_QUALITY, q] + extra + sflags)
                    assert ok
                    out.append(((h, w, q, label, sname), buf.tobytes()))
    return out


def test_oracle_matches_cv2_on_fresh_encodes():
    """Streams the fixtures do not hold: odd sizes, optimised Huffman tables (DHTs that differ from stream to
    stream), restart intervals, 4:4:4 -- encoded and decoded by the installed cv2 on the spot."""
    cv2 = pytest.importorskip('cv2')
    n = 0
    for (tag, data) in fresh_encodes():
        want = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
        assert np.array_equal(oracle.decode_jpeg(data), want), tag
        n += 1
    assert n >= 36

