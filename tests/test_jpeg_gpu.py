"""Ingest row on the GPU: melf_decode_jpeg against the CPU oracle and against
what cv2.imread returned in the build container; melf_read_jpeg_batch against the reference's
golden readings.  Everything goes through the C-ABI."""
import collections
import hashlib
import os

import numpy as np
import pytest

import oracle
from helpers import GOLDEN, assert_results_equal, format_result_line, golden_set
from test_jpeg_oracle import load_streams

pytestmark = pytest.mark.gpu


def decode_on_gpu(eng, streams, h, w):
    import torch
    from meterelf_b200 import _engine
    (blob, offsets) = _engine.pack_streams(streams)
    out = torch.zeros((len(streams), h, w, 3), dtype=torch.uint8, device='cuda:0')
    eng.decode_jpeg_device(blob, offsets, h, w, out.data_ptr(), w * 3, h * w * 3,
                           torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def batches(streams):
    """Streams grouped into batches of one geometry (size + subsampling), as the entry point wants."""
    from meterelf_b200 import _engine
    groups = collections.OrderedDict()
    for s in streams:
        info = _engine.jpeg_info(s[1])
        groups.setdefault(tuple(sorted(info.items())), []).append(s)
    return list(groups.values())


def test_committed_streams_decode_like_cv2_and_the_oracle():
    eng = golden_set(1).engine()
    streams = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    n = 0
    for group in batches(streams):
        (h, w) = group[0][2]
        got = decode_on_gpu(eng, [s[1] for s in group], h, w)
        for (img, (name, data, shape, sha, _)) in zip(got, group):
            want = oracle.decode_jpeg(data)
            assert np.array_equal(img, want), (name, np.argwhere(img != want)[:4])
            assert hashlib.sha256(img.tobytes()).hexdigest() == sha, name
            n += 1
    assert n == len(streams) >= 50


def test_every_sample_file_decodes_like_cv2():
    path = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    if not os.path.exists(path):
        pytest.skip('full JPEG set not present')
    eng = golden_set(1).engine()
    streams = load_streams(path)
    assert len(streams) == 304
    for group in batches(streams):
        (h, w) = group[0][2]
        got = decode_on_gpu(eng, [s[1] for s in group], h, w)
        for (img, (name, _, _, sha, _)) in zip(got, group):
            assert hashlib.sha256(img.tobytes()).hexdigest() == sha, name


def test_decode_into_strided_frames_and_repeated_calls():
    import torch
    from meterelf_b200 import _engine
    eng = golden_set(1).engine()
    streams = [s for s in load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz')) if s[2] == (17, 33)]
    group = batches(streams)[0]
    (h, w) = (17, 33)
    (blob, offsets) = _engine.pack_streams([s[1] for s in group])
    (rs, fs) = (w * 3 + 13, (w * 3 + 13) * h + 7)
    for _ in range(2):
        buf = torch.full((len(group) * fs,), 0xA5, dtype=torch.uint8, device='cuda:0')
        eng.decode_jpeg_device(blob, offsets, h, w, buf.data_ptr(), rs, fs)
        torch.cuda.synchronize()
        host = buf.cpu().numpy()
        for (i, s) in enumerate(group):
            frame = host[i * fs:i * fs + rs * h].reshape(h, rs)
            assert np.array_equal(frame[:, :w * 3].reshape(h, w, 3), oracle.decode_jpeg(s[1])), s[0]
            assert (frame[:, w * 3:] == 0xA5).all()          # nothing written outside the rows


def test_unsupported_and_damaged_streams_fail_loudly():
    from meterelf_b200 import _engine
    eng = golden_set(1).engine()
    (name, data, shape, _, _) = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))[0]
    prog = bytearray(data)
    prog[prog.find(b'\xff\xc0') + 1] = 0xC2
    for bad in (bytes(prog), b'junk', data[:300]):
        with pytest.raises(_engine.EngineError):
            eng.read_jpeg_streams([data, bad])
    with pytest.raises(_engine.EngineError):                 # size mismatch inside one batch
        decode_on_gpu(eng, [data], shape[0] + 1, shape[1])


def test_streams_cut_short_decode_like_libjpeg():
    """Truncated entropy data (jdhuff.c insufficient_data) through melf_decode_jpeg: equal to the oracle
    for every committed stream, restart intervals included, and to cv2.imread's vectors where they exist."""
    import json
    eng = golden_set(1).engine()
    streams = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    with open(os.path.join(GOLDEN, 'jpeg_truncated.json')) as fp:
        vectors = {(c['name'], c['keep']): c['sha256'] for c in json.load(fp)}
    (n, n_vec) = (0, 0)
    for frac in (0.35, 0.5, 0.77, 0.93, 0.999):
        cut = []
        for s in streams:
            part = s[1][:int(len(s[1]) * frac)]
            try:
                oracle.decode_jpeg(part)
            except oracle.JpegError:              # cut inside the headers
                continue
            cut.append((s[0], part, s[2], None, s[4]))
        for group in batches(cut):
            (h, w) = group[0][2]
            got = decode_on_gpu(eng, [s[1] for s in group], h, w)
            for (img, s) in zip(got, group):
                want = oracle.decode_jpeg(s[1])
                assert np.array_equal(img, want), (s[0], frac, np.argwhere(img != want)[:3])
                sha = vectors.get((s[0], len(s[1])))
                if sha:
                    assert hashlib.sha256(img.tobytes()).hexdigest() == sha, (s[0], frac)
                    n_vec += 1
                n += 1
    assert n >= 150 and n_vec >= 60


def test_intact_and_cut_streams_in_one_warp():
    """The lean Huffman loop only FLAGS a stream that read past its end; the careful redo pass behind it clears and
    decodes the flagged streams again and must leave their neighbours in the warp alone.  One batch, intact and
    cut streams alternating (several cut points, so that flagged and unflagged lanes share every warp), twice in a
    row on the same buffers."""
    eng = golden_set(1).engine()
    streams = load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    base = max(batches([s for s in streams if s[2] == (640, 480)]), key=len)      # one geometry, most streams
    assert len(base) >= 3
    mixed = []
    for (k, s) in enumerate((base * 16)[:48]):
        frac = (1.0, 0.5, 1.0, 0.93, 0.77, 1.0, 0.35, 0.999)[k % 8]
        mixed.append(s[1] if frac == 1.0 else s[1][:int(len(s[1]) * frac)])
    want = [oracle.decode_jpeg(m) for m in mixed]
    for _ in range(2):
        got = decode_on_gpu(eng, mixed, 640, 480)
        for (k, (img, ref)) in enumerate(zip(got, want)):
            assert np.array_equal(img, ref), (k, len(mixed[k]), np.argwhere(img != ref)[:3])


def test_fresh_encodes_mixed_tables_restarts_and_cuts():
    """Batches the committed streams do not offer: DHTs that differ from stream to stream (one table set per lane
    instead of one per CTA), with and without restart intervals (careful loop / lean loop + redo pass), intact and
    cut streams side by side.  Intact streams against cv2.imdecode, cut ones against the oracle."""
    cv2 = pytest.importorskip('cv2')
    from test_jpeg_oracle import fresh_encodes
    eng = golden_set(1).engine()
    enc = fresh_encodes(seed=5, sizes=((120, 200), (481, 643)))
    for sampling in ('420', '444'):
        for with_restarts in (False, True):
            for (h, w) in ((120, 200), (481, 643)):
                group = [(t, d) for (t, d) in enc if t[:2] == (h, w) and t[4] == sampling and
                         (with_restarts or t[3] != 'rst')]
                if not group:
                    continue
                streams = []
                for (k, (_t, d)) in enumerate(group * 3):
                    frac = (1.0, 0.6, 1.0, 0.9)[k % 4]
                    streams.append(d if frac == 1.0 else d[:int(len(d) * frac)])
                got = decode_on_gpu(eng, streams, h, w)
                for (k, (img, d)) in enumerate(zip(got, streams)):
                    if len(d) == len(group[k % len(group)][1]):
                        want = cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR)
                    else:
                        want = oracle.decode_jpeg(d)
                    assert np.array_equal(img, want), (group[k % len(group)][0], len(d))


def test_readings_from_streams_cut_short_match_the_oracle():
    """melf_read_jpeg_batch on truncated streams: the rectangle-only decode (kept / skipped MCU runs, early
    stop) has to leave the same grey MCUs behind as the full decode, so the readings -- or the errors --
    are the oracle's on the oracle's decode of the same bytes."""
    gs = golden_set(1)
    eng = gs.engine(cropped_input=False)
    streams = [s for s in load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz')) if s[4] == 1 and s[2] == (640, 480)]
    assert streams
    (x0, y0, x1, y1) = gs.meter_rect
    cut = []
    for s in streams:
        for frac in (0.3, 0.45, 0.55, 0.62, 0.7, 0.999):
            cut.append(s[1][:int(len(s[1]) * frac)])
    got = eng.read_jpeg_streams(cut)
    crops = np.stack([oracle.decode_jpeg(d)[y0:y1, x0:x1] for d in cut])
    want = oracle.read_batch(gs.oracle_params(), np.ascontiguousarray(crops), n_threads=4)
    statuses = set()
    for (i, (g, w)) in enumerate(zip(got, want)):
        assert_results_equal(g, w, len(gs.names), i)
        statuses.add(int(w['status']))
    assert len(statuses) >= 2           # some cuts lose the dials, the late ones do not


@pytest.mark.parametrize('set_no', [1, 2])
def test_readings_from_jpeg_bytes_match_the_golden_lines(set_no):
    path = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    streams = load_streams(path if os.path.exists(path) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
    streams = [s for s in streams if s[4] == set_no]
    gs = golden_set(set_no)
    eng = gs.engine(cropped_input=False)
    n_lines = 0
    for group in batches(streams):
        got = eng.read_jpeg_streams([s[1] for s in group])
        # the same frames decoded by the oracle, cropped as _image.py:46-55 does, read by the oracle
        (x0, y0, x1, y1) = gs.meter_rect
        crops = np.stack([oracle.decode_jpeg(s[1])[y0:y1, x0:x1] for s in group[:12]])
        want = oracle.read_batch(gs.oracle_params(), np.ascontiguousarray(crops), n_threads=4)
        for (g, w, s) in zip(got, want, group):
            assert_results_equal(g, w, len(gs.names), s[0])
        for (g, s) in zip(got, group):
            exp = gs.expected[s[0]]
            line = format_result_line(g, gs.names)
            if exp['status'] == 0:
                assert line == exp['line'], s[0]
                n_lines += 1
            else:
                assert line.split(' (match val')[0] == exp['line'].split(' (match val')[0], s[0]
    assert n_lines >= 2


def test_colour_space_variants_ycbcr_on_the_gpu_rgb_back_to_the_host(tmp_path):
    """streams libjpeg reads as YCbCr (JFIF, Adobe transform 1/2, any ids but R,G,B)
    decode on the GPU like cv2 did; streams it reads as RGB (Adobe transform 0 / ids R,G,B without
    JFIF) are refused by melf_decode_jpeg and reach the reading path through cv2.imread."""
    import json
    from helpers import jpeg_colour_space_bases, jpeg_colour_space_variants, write_params_file
    from meterelf_b200 import _engine
    with open(os.path.join(GOLDEN, 'jpeg_colorspace.json')) as fp:
        vectors = {(v['name'], v['variant']): v for v in json.load(fp)}
    eng = golden_set(1).engine()
    for (name, data) in jpeg_colour_space_bases():
        variants = jpeg_colour_space_variants(data)
        ycc = [(key, v) for (key, v, is_rgb) in variants if not is_rgb]
        (h, w) = vectors[(name, 'jfif')]['shape']
        got = decode_on_gpu(eng, [v for (_k, v) in ycc], h, w)
        for (img, (key, _v)) in zip(got, ycc):
            assert hashlib.sha256(img.tobytes()).hexdigest() == vectors[(name, key)]['sha256'], (name, key)
        for (key, v, is_rgb) in variants:
            if is_rgb:
                with pytest.raises(_engine.EngineError) as exc:
                    decode_on_gpu(eng, [v], h, w)
                assert exc.value.code == _engine.ERR_UNSUPPORTED
    # through the Python API: an RGB-labelled sample file gives what cv2.imread gives
    import cv2
    import meterelf_b200
    gs = golden_set(1)
    params_file = write_params_file(gs, tmp_path)
    (name, data) = jpeg_colour_space_bases()[0]
    paths = []
    for (key, v, is_rgb) in jpeg_colour_space_variants(data):
        if key in ('jfif', 'adobe0', 'ids_rgb'):
            paths.append(str(tmp_path / (key + '.jpg')))
            with open(paths[-1], 'wb') as fp:
                fp.write(v)
    got = list(meterelf_b200.get_meter_values(params_file, paths))
    (x0, y0, x1, y1) = gs.meter_rect
    crops = np.stack([np.ascontiguousarray(cv2.imread(p)[y0:y1, x0:x1]) for p in paths])
    want = oracle.read_batch(gs.oracle_params(), crops)
    for (g, wnt, p) in zip(got, want, paths):
        if int(wnt['status']) == 0:
            assert g.error is None and g.value == float(wnt['value']), p
        else:
            assert g.error is not None and g.value is None, p
            assert ('UNKNOWN ' + g.error.get_message()) == format_result_line(wnt, gs.names), p
