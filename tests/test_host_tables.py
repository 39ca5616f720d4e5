"""Host-side static tables of the CUDA library, checked on the CPU (no GPU needed).

* the tensor-core TM_CCOEFF formulation (Toeplitz operand image + strip layout + M-tile
  plan, modelled in numpy by ``kernel_model.match_tc_model``) reproduces the direct
  correlation sum for the sample geometry and for ragged sizes;
* the per-dial tables (``melf_host_dial_tables``) agree with the masks they come from.
"""
import ctypes as C

import numpy as np
import pytest

import kernel_model
from helpers import golden_set
from meterelf_b200 import _engine


def _lib():
    return _engine.load_library()


def _toeplitz(tmpl):
    lib = _lib()
    lib.melf_host_match_tc_toeplitz.restype = C.c_longlong
    lib.melf_host_match_tc_toeplitz.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_longlong]
    (th, tw) = tmpl.shape
    need = lib.melf_host_match_tc_toeplitz(tmpl.ctypes.data, tw, th, None, 0)
    assert need == ((th + 1) // 2) * 32768
    buf = np.zeros(need, np.uint8)
    assert lib.melf_host_match_tc_toeplitz(tmpl.ctypes.data, tw, th, buf.ctypes.data, need) == need
    return buf.tobytes()


def _plan(res_w, res_h, tw, th):
    lib = _lib()
    plan = (C.c_int32 * 4)()
    assert lib.melf_host_match_tc_plan(res_w, res_h, tw, th, plan) == 0
    return tuple(plan)


def _direct(light, tmpl):
    (th, tw) = tmpl.shape
    (h, w) = light.shape
    win = np.lib.stride_tricks.sliding_window_view(light.astype(np.int64), (th, tw))
    return np.einsum('yxij,ij->yx', win, tmpl.astype(np.int64))


def test_match_tc_model_sample_geometry():
    gs = golden_set(1)
    crop = gs.subset()['crops'][3]
    light = (crop.max(axis=2).astype(np.int32) + crop.min(axis=2) + 1) // 2     # any u8 plane will do
    light = light.astype(np.uint8)
    tmpl = np.ascontiguousarray(gs.template)
    (th, tw) = tmpl.shape
    plan = _plan(light.shape[1] - tw + 1, light.shape[0] - th + 1, tw, th)
    assert plan == (1, 1, 132, 60)
    got = kernel_model.match_tc_model(light, _toeplitz(tmpl), tw, th, plan)
    assert np.array_equal(got, _direct(light, tmpl))


@pytest.mark.parametrize('crop_hw,tmpl_hw', [((135, 220), (119, 188)), ((40, 100), (7, 30)),
                                              ((300, 200), (20, 60)), ((150, 321), (12, 193)),
                                              ((128 + 8, 70), (9, 5))])
def test_match_tc_model_ragged_sizes(crop_hw, tmpl_hw):
    rng = np.random.default_rng(crop_hw[0] * 1000 + tmpl_hw[1])
    light = rng.integers(0, 256, size=crop_hw, dtype=np.uint8)
    tmpl = rng.integers(0, 256, size=tmpl_hw, dtype=np.uint8)
    (th, tw) = tmpl_hw
    plan = _plan(crop_hw[1] - tw + 1, crop_hw[0] - th + 1, tw, th)
    got = kernel_model.match_tc_model(light, _toeplitz(tmpl), tw, th, plan)
    assert np.array_equal(got, _direct(light, tmpl))


def test_match_tc_unsupported_templates_report_it():
    lib = _lib()
    lib.melf_host_match_tc_toeplitz.restype = C.c_longlong
    lib.melf_host_match_tc_toeplitz.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_longlong]
    assert lib.melf_host_match_tc_toeplitz(None, 194, 100, None, 0) == -1     # K = tw + 63 > 256
    assert lib.melf_host_match_tc_toeplitz(None, 100, 129, None, 0) == -1


@pytest.mark.parametrize('set_no', [1, 2])
def test_dial_tables_match_masks(set_no):
    gs = golden_set(set_no)
    lib = _lib()
    (cp, keep) = gs.cparams()
    for d in range(len(gs.names)):
        geom = (C.c_int32 * 6)()
        words = (C.c_uint64 * (2 + 2 * 64))()
        terms = (C.c_double * 128)()
        n_ring = C.c_int32()
        cap = 4096
        pos = (C.c_uint16 * cap)()
        ang = (C.c_double * cap)()
        sqd = (C.c_double * cap)()
        rc = lib.melf_host_dial_tables(C.byref(cp), d, geom, words, terms, C.byref(n_ring), pos, ang, sqd, cap)
        assert rc == 0
        (wx, wy, ww, wh, px, py) = list(geom)
        mask = gs.dials[d]['mask']
        ring = gs.dials[d]['circle_mask']
        (cx, cy) = gs.dials[d]['center']
        assert (px, py) == (int(cx) - 2, int(cy) - 2)
        disc = np.array(list(words)[2:66], np.uint64)
        ringw = np.array(list(words)[66:130], np.uint64)
        n_disc = sum(bin(int(w)).count('1') for w in disc)
        assert n_disc == int(np.count_nonzero(mask))
        assert sum(bin(int(w)).count('1') for w in ringw) == int(np.count_nonzero(ring))
        # every ring pixel with a defined angle is in the sorted table, ascending by (angle, sqd)
        n = n_ring.value
        assert 0 < n <= int(np.count_nonzero(ring))
        keys = [(ang[i], sqd[i]) for i in range(n)]
        assert keys == sorted(keys)
        for i in range(0, n, 7):
            (r, c) = (pos[i] >> 8, pos[i] & 255)
            (x, y) = (wx + c, wy + r)
            assert ring[y, x]
            a = kernel_model_angle(x - cx, y - cy)
            assert a == ang[i]
            assert sqd[i] == (x - cx) ** 2 + (y - cy) ** 2


def kernel_model_angle(x, y):
    import math
    if y == 0:
        return 0.25 if x > 0 else 0.75
    return (-(math.atan(x / y) / (2 * math.pi)) + (0.5 if y > 0 else 0.0)) % 1.0


def test_jpeg_ingest_header_helpers():
    """Which files the API hands to the ingest kernels as bytes (meterelf_b200/_image.py)."""
    import os
    import struct
    from helpers import GOLDEN
    from test_jpeg_oracle import load_streams
    from meterelf_b200 import _image
    streams = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    (name, data, shape, _sha, _set) = streams[0]
    geometry = _image.gpu_ingest_geometry(data)
    assert geometry is not None and geometry[:3] == (shape[0], shape[1], 3)
    assert _image.exif_orientation(data) == 1
    assert _image.gpu_ingest_geometry(b'\x89PNG\r\n\x1a\n' + b'\0' * 64) is None
    assert _image.gpu_ingest_geometry(data[:100]) is None            # headers cut short
    assert _image.gpu_ingest_geometry(data[:len(data) // 2]) is None  # entropy data cut short: no EOI
    assert _image.gpu_ingest_geometry(data + b'\0' * 20) == geometry  # padding after EOI is fine

    def with_exif(orientation, order):
        fmt = '<' if order == b'II' else '>'
        tiff = order + struct.pack(fmt + 'HI', 42, 8) + struct.pack(fmt + 'H', 1) + \
            struct.pack(fmt + 'HHIHH', 0x0112, 3, 1, orientation, 0) + struct.pack(fmt + 'I', 0)
        seg = b'Exif\0\0' + tiff
        return data[:2] + b'\xff\xe1' + struct.pack('>H', len(seg) + 2) + seg + data[2:]

    for order in (b'II', b'MM'):
        assert _image.exif_orientation(with_exif(6, order)) == 6
        assert _image.gpu_ingest_geometry(with_exif(6, order)) is None     # cv2.imread would rotate
        assert _image.gpu_ingest_geometry(with_exif(1, order)) == geometry


def test_debug_overlay_helpers(monkeypatch):
    """PNG writer, tap-window unpacking and the on/off switch of meterelf_b200/_overlay.py."""
    import struct
    import zlib
    from meterelf_b200 import _debug, _overlay
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, size=(13, 21, 3), dtype=np.uint8)
    png = _overlay.encode_png(img)
    assert png[:8] == b'\x89PNG\r\n\x1a\n'
    (w, h, depth, ctype) = struct.unpack('>IIBB', png[16:26])
    assert (w, h, depth, ctype) == (21, 13, 8, 2)
    idat_len = struct.unpack('>I', png[33:37])[0]
    raw = np.frombuffer(zlib.decompress(png[41:41 + idat_len]), np.uint8).reshape(13, 1 + 63)
    assert not raw[:, 0].any() and np.array_equal(raw[:, 1:].reshape(13, 21, 3), img[:, :, ::-1])
    try:
        import cv2
        assert np.array_equal(cv2.imdecode(np.frombuffer(png, np.uint8), cv2.IMREAD_COLOR), img)
    except ImportError:
        pass
    # scale_image = cv2.resize's fixed-point bilinear filter (_utils.py:57-61), hls_to_bgr_full =
    # cv2.cvtColor(HLS2BGR_FULL) in the operation order of OpenCV's vectorised code
    big = _overlay.scale_image(img, 4)
    assert big.shape == (52, 84, 3)
    try:
        import cv2
        cv2.ipp.setUseIPP(False)
        for (h, w) in ((13, 21), (1, 1), (2, 33), (119, 188)):
            src = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
            for scale in (2, 4):
                assert np.array_equal(_overlay.resize_linear(src, scale), cv2.resize(src, (w * scale, h * scale)))
        (L, S) = np.meshgrid(np.arange(256), np.arange(256), indexing='ij')
        differing = 0
        for hue in range(0, 256):
            hls = np.stack([np.full_like(L, hue), L, S], axis=-1).astype(np.uint8)
            want = cv2.cvtColor(hls, cv2.COLOR_HLS2BGR_FULL)
            got = _overlay.hls_to_bgr_full(hls)
            bad = np.any(got != want, axis=-1)
            differing += int(bad.sum())
            assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
        assert differing <= 8, differing      # 4 of 2^24 colours: float32 near-ties (rint of x.49998 / x.50002)
    except ImportError:
        pass
    words = np.zeros(64, np.uint64)
    words[0] = (1 << 3) | (1 << 63)
    words[63] = 1
    mask = _overlay.window_to_mask(words, -2, 5, (70, 80))
    assert mask[5, 1] and mask[5, 61] and mask.sum() == 2          # (68, -2) falls outside of the image
    monkeypatch.setenv('MELF_DEBUG_DIR', '/tmp/x')
    monkeypatch.setattr(_debug, 'DEBUG', set())
    assert _overlay.overlay_dir() is None
    monkeypatch.setattr(_debug, 'DEBUG', {'1'})
    assert _overlay.overlay_dir() == '/tmp/x'
    monkeypatch.delenv('MELF_DEBUG_DIR')
    assert _overlay.overlay_dir() is None


def test_batch_file_reader_and_header_probe(tmp_path):
    """melf_file_sizes / melf_read_files / melf_jpeg_probe_batch (the per-file preamble of
    get_meter_values, batched behind the C ABI) against the one-stream Python forms."""
    import os
    from helpers import GOLDEN, jpeg_colour_space_bases, jpeg_colour_space_variants
    from test_jpeg_oracle import load_streams
    from meterelf_b200 import _engine, _image
    streams = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    cases = [(name, data) for (name, data, _s, _h, _n) in streams]
    (name0, data0) = cases[0]
    cases.append(('cut.jpg', data0[:len(data0) // 2]))                  # no end-of-image marker
    cases.append(('head.jpg', data0[:100]))                             # headers cut short
    cases.append(('not.jpg', b'\x89PNG\r\n\x1a\n' + b'\0' * 64))
    cases.append(('empty.jpg', b''))
    prog = bytearray(data0)
    prog[prog.find(b'\xff\xc0') + 1] = 0xC2
    cases.append(('progressive.jpg', bytes(prog)))
    for (key, variant, _rgb) in jpeg_colour_space_variants(jpeg_colour_space_bases()[0][1]):
        cases.append(('cs-%s.jpg' % key, variant))
    paths = []
    for (name, data) in cases:
        paths.append(str(tmp_path / name))
        with open(paths[-1], 'wb') as fp:
            fp.write(data)
    paths.append(str(tmp_path / 'missing.jpg'))
    (blob, offsets, ok) = _engine.read_files(paths)
    assert list(ok) == [True] * len(cases) + [False]
    for (i, (name, data)) in enumerate(cases):
        assert blob[int(offsets[i]):int(offsets[i + 1])].tobytes() == data, name
    assert offsets[-1] == offsets[-2]                                   # the missing file: an empty slice
    info = _engine.jpeg_probe_batch(blob, offsets)
    assert info.shape == (len(paths), 8)
    n_gpu = 0
    for (i, (name, data)) in enumerate(cases):
        want = _image.gpu_ingest_geometry(data) if data else None
        row = info[i]
        takes = bool(row[7] == 0 and row[5] == 1 and row[6] == 1)
        assert takes == (want is not None), (name, row, want)
        if want is not None:
            assert tuple(int(v) for v in row[:5]) == want, name
            n_gpu += 1
    assert n_gpu >= 50 and info[-1, 7] != 0
    assert len(_engine.read_files([])[1]) == 1 and _engine.jpeg_probe_batch(blob, offsets[:1]).shape == (0, 8)
