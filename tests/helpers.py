"""Shared test plumbing: golden fixtures -> oracle params / engine params."""
import functools
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, 'golden')


class GoldenSet:
    def __init__(self, set_no):
        self.set_no = set_no
        P = np.load(os.path.join(GOLDEN, 'set%d_params.npz' % set_no))
        self.template = P['template']
        (self.tmpl_h, self.tmpl_w) = self.template.shape
        self.names = [str(s) for s in P['names']]
        self.meter_rect = tuple(int(v) for v in P['meter_rect'])
        self.hue_shift = int(P['hue_shift'])
        self.match_threshold = float(P['match_threshold'])
        self.centers = P['centers']
        self.diameters = P['diameters']
        self.dist_from_center = P['dist_from_center']
        self.circle_thickness = P['circle_thickness']
        masks = np.unpackbits(P['ref_mask'], axis=-1)[..., :self.tmpl_w] * np.uint8(255)
        rings = np.unpackbits(P['ref_circle_mask'], axis=-1)[..., :self.tmpl_w] * np.uint8(255)
        self.dials = [dict(
            name=self.names[i], center=(float(P['centers'][i][0]), float(P['centers'][i][1])),
            color_range=tuple(int(c) for c in P['color_range'][i]),
            negative_momentum=bool(P['negative_momentum'][i]),
            angle_of_zero=float(P['angle_of_zero'][i]),
            mask=np.ascontiguousarray(masks[i]), circle_mask=np.ascontiguousarray(rings[i]))
            for i in range(len(self.names))]
        expected = os.path.join(GOLDEN, 'set%d_expected.json' % set_no)
        self.expected = {}                  # set 3 (discs cut by the rect's frame) has no sample frames
        if os.path.exists(expected):
            with open(expected) as fp:
                self.expected = json.load(fp)['frames']

    def oracle_params(self, threshold=None, dials=None):
        import oracle
        return oracle.OracleParams(
            self.template, self.hue_shift,
            self.match_threshold if threshold is None else threshold,
            self.dials if dials is None else dials)

    def engine(self, cropped_input=True, threshold=None, dials=None, device=0):
        from meterelf_b200 import _engine
        rect = (0, 0, 2 ** 30, 2 ** 30) if cropped_input else self.meter_rect
        (cp, keep) = _engine.build_cparams(
            self.template, rect, self.hue_shift,
            self.match_threshold if threshold is None else threshold,
            self.dials if dials is None else dials)
        return _engine.Engine(cp, keep, self.names, device)

    def cparams(self, cropped_input=True):
        from meterelf_b200 import _engine
        rect = (0, 0, 2 ** 30, 2 ** 30) if cropped_input else self.meter_rect
        return _engine.build_cparams(self.template, rect, self.hue_shift,
                                     self.match_threshold, self.dials)

    def subset(self):
        return np.load(os.path.join(GOLDEN, 'set%d_subset.npz' % self.set_no))

    def frames(self):
        """(names, crops): every sample frame when the git-ignored full set is present
        (build container / gpurun snapshot), else the committed subset."""
        full = os.path.join(GOLDEN, '_full', 'set%d_crops.npz' % self.set_no)
        data = np.load(full) if os.path.exists(full) else self.subset()
        return ([str(s) for s in data['names']], data['crops'])


@functools.lru_cache(maxsize=None)
def golden_set(set_no):
    return GoldenSet(set_no)


def format_result_line(rec, names):
    """What the CLI prints after '<file>: ' for a result record (_main.py:16-22)."""
    st = int(rec['status'])
    if st == 0:
        v = float(rec['value'])
        return '{:07.3f}'.format(v) if v else ''
    if st == 1:
        return 'UNKNOWN Dials not found (match val = {})'.format(float(rec['max_val']))
    if st == 2:
        return 'UNKNOWN Cannot find needle contours of a dial (dial = {})'.format(
            names[int(rec['failed_dial'])])
    bad = [n for (i, n) in enumerate(names) if (int(rec['unreadable_mask']) >> i) & 1]
    return 'UNKNOWN Cannot determine angle of a dial (unreadable dials = {})'.format(
        ', '.join(bad))


def assert_results_equal(got, want, n_dials, context=''):
    """Field-by-field equality of result records; float64 fields bit-exact (NaN == NaN)."""
    for k in ('status', 'match_x', 'match_y', 'failed_dial', 'unreadable_mask'):
        assert int(got[k]) == int(want[k]), (context, k, got, want)
    assert np.float32(got['max_val']).tobytes() == np.float32(want['max_val']).tobytes(), (
        context, 'max_val', got['max_val'], want['max_val'])
    g = np.asarray(got['dial_pos'][:n_dials], np.float64)
    w = np.asarray(want['dial_pos'][:n_dials], np.float64)
    assert g.tobytes() == w.tobytes() or (
        np.array_equal(np.isnan(g), np.isnan(w)) and np.array_equal(g[~np.isnan(g)], w[~np.isnan(w)])), (
        context, 'dial_pos', g, w)
    gv, wv = float(got['value']), float(want['value'])
    assert (gv != gv and wv != wv) or gv == wv, (context, 'value', gv, wv)


def window_words(mask, win_x0, win_y0):
    """u8 mask (h, w) -> 64 row words of the dial window (bit c = column win_x0 + c)."""
    (h, w) = mask.shape
    out = np.zeros(64, np.uint64)
    for r in range(64):
        y = win_y0 + r
        if not 0 <= y < h:
            continue
        word = 0
        for c in range(64):
            x = win_x0 + c
            if 0 <= x < w and mask[y, x]:
                word |= 1 << c
        out[r] = word
    return out


def synthetic_blob_frames(rng, n, h, w, fg=(40, 60, 200), bg=(120, 130, 125), cell=(3, 9),
                          density=(0.2, 0.6), noise=6):
    """Random two-colour frames: blobs of a needle-like colour on a grey background, so
    that the per-dial inRange masks become arbitrary shapes (holes, islands, diagonal
    links, specks).  Exercises the branches the sample images never reach."""
    frames = np.empty((n, h, w, 3), np.uint8)
    for i in range(n):
        c = int(rng.integers(cell[0], cell[1] + 1))
        gh, gw = h // c + 2, w // c + 2
        coarse = rng.random((gh, gw)) < rng.uniform(*density)
        mask = np.kron(coarse, np.ones((c, c), bool))[:h, :w]
        speck = rng.random((h, w)) < 0.03
        mask = mask ^ speck
        img = np.where(mask[..., None], np.array(fg, np.int16), np.array(bg, np.int16))
        img = img + rng.integers(-noise, noise + 1, size=(h, w, 3))
        frames[i] = np.clip(img, 0, 255).astype(np.uint8)
    return frames


# (case key, params set, seed, frames): the batches tests/golden/make_golden_synthetic.py ran the
# unmodified reference on (tests/golden/synthetic_expected.json)
# seeds >= 5: large dense blobs (more dials readable, more components per disc)
SYNTHETIC_CASES = [('set2_seed0', 2, 0, 160), ('set2_seed1', 2, 1, 160), ('set1_seed2', 1, 2, 48),
                   ('set3_seed3', 3, 3, 160), ('set3_seed4', 3, 4, 160),
                   ('set2_seed5', 2, 5, 160), ('set3_seed6', 3, 6, 160), ('set1_seed7', 1, 7, 48)]


def synthetic_batch(seed, n, h, w):
    """Blob frames of one synthetic case plus two degenerate ones: flat grey (every TM value 0 ->
    match at (0, 0)) and flat needle colour (everything in range)."""
    kwargs = dict(cell=(4, 14), density=(0.45, 0.9)) if seed >= 5 else {}
    frames = synthetic_blob_frames(np.random.default_rng(seed), n, h, w, **kwargs)
    frames[0] = 128
    frames[1, :, :] = (40, 60, 200)
    return frames


@functools.lru_cache(maxsize=1)
def synthetic_expected():
    with open(os.path.join(GOLDEN, 'synthetic_expected.json')) as fp:
        return json.load(fp)


NEAR_TIE_TOLERANCE = 256.0    # |exact TM_CCOEFF - cv2's float32 DFT value|, DESIGN.md section 2


def check_against_reference_record(gs, oracle_params, frame, res, taps, exp, context):
    """One result record (oracle or GPU; `taps` = the oracle's taps or None) against what the
    unmodified reference computed on the same frame (a record of synthetic_expected.json).

    Match location: the exact integer TM_CCOEFF arg-max equals cv2's unless the best exact score
    leads another position by less than cv2's float32-DFT error.  Near-tie rule: when the
    locations differ, the exact score at cv2's location must lie within NEAR_TIE_TOLERANCE of the
    exact maximum (and cv2's max_val within the tolerance of it); the dial stage is then checked
    by reading the dials rect at cv2's location with the oracle.  Returns 1 for a near-tie frame.
    """
    import oracle
    n_dials = len(gs.names)
    assert abs(float(res['max_val']) - exp['cv2_max_val']) <= NEAR_TIE_TOLERANCE, context
    near_tie = (int(res['match_x']), int(res['match_y'])) != (exp['match_x'], exp['match_y'])
    if near_tie:
        hls = oracle.convert_to_hls(frame, gs.hue_shift)
        (_mx, _my, mv, tm) = oracle.match_template(np.ascontiguousarray(hls[:, :, 1]), gs.template)
        at_cv2 = float(tm[exp['match_y'], exp['match_x']])
        assert float(mv) - at_cv2 <= NEAR_TIE_TOLERANCE, (context, float(mv), at_cv2)
        rect = np.ascontiguousarray(hls[exp['match_y']:exp['match_y'] + gs.tmpl_h,
                                        exp['match_x']:exp['match_x'] + gs.tmpl_w])
        (res, taps) = oracle.read_dials_hls(oracle_params, rect, want_taps=True)
    assert int(res['status']) == exp['status'], (context, int(res['status']), exp)
    if exp['status'] == 0:
        want = [float.fromhex(v) for v in exp['dial_pos']]
        assert [float(v) for v in res['dial_pos'][:n_dials]] == want, (context, 'dial_pos')
        assert float(res['value']) == float.fromhex(exp['value']), (context, 'value')
    elif exp['status'] == 2:
        assert int(res['failed_dial']) == exp['failed_dial'], context
    else:
        assert int(res['unreadable_mask']) == exp['unreadable_mask'], context
    assert format_result_line(res, gs.names).replace('UNKNOWN ', '') == exp.get('message', format_result_line(res, gs.names)), context
    if taps is not None:
        for d in range(n_dials):
            if exp['n_needle'][d] is None:           # the reference raised before this dial
                continue
            assert int(taps.n_needle_points[d]) == exp['n_needle'][d], (context, d, 'n_needle')
            assert float(taps.contour_area[d]) == exp['area'][d], (context, d, 'area')
    return 1 if near_tie else 0


JPEG_COLOUR_SPACE_BASES = ['20180814120000-e50-snapshot.jpg', 'synthetic-48x64-444-rst0-q75',
                           'synthetic-188x119-422-rst3-q100', 'synthetic-33x17-420-rst0-q75']


def jpeg_colour_space_bases():
    """[(name, bytes)]: the committed streams the colour space variants are made from."""
    d = np.load(os.path.join(GOLDEN, 'jpeg_subset.npz'))
    names = [str(n) for n in d['names']]
    blob = d['blob'].tobytes()
    off = d['offsets']
    return [(n, blob[off[names.index(n)]:off[names.index(n) + 1]]) for n in JPEG_COLOUR_SPACE_BASES]


# frames whose debug pictures tests/golden/make_golden_overlay.py captured from the reference:
# sample crops by file name, synthetic blob frames by (case of SYNTHETIC_CASES, index)
OVERLAY_CASES = [
    dict(key='set1:20180814215230-01-e136.jpg', set=1, name='20180814215230-01-e136.jpg', threshold=None),
    dict(key='set2:20181001_113527-01-e14.jpg', set=2, name='20181001_113527-01-e14.jpg', threshold=None),
] + [dict(key='%s[%d]' % (case, i), set=set_no, synthetic=(case, i), threshold=-1e30)
     for (case, set_no, picks) in (('set2_seed5', 2, (2, 3, 5, 8, 13, 21, 24, 30, 34)), ('set3_seed6', 3, (2, 4, 7, 11, 16, 31)),
                                   ('set1_seed7', 1, (3, 6, 9, 10, 14, 20)))
     for i in picks]


def overlay_case_frame(case):
    """The meter_rect crop of one OVERLAY_CASES entry."""
    gs = golden_set(case['set'])
    if 'name' in case:
        sub = gs.subset()
        names = [str(n) for n in sub['names']]
        return np.ascontiguousarray(sub['crops'][names.index(case['name'])])
    (key, index) = case['synthetic']
    (seed, n) = next((c[2], c[3]) for c in SYNTHETIC_CASES if c[0] == key)
    (x0, y0, x1, y1) = gs.meter_rect
    return np.ascontiguousarray(synthetic_batch(seed, n, y1 - y0, x1 - x0)[index])


def jpeg_colour_space_variants(data):
    """[(key, bytes, libjpeg reads it as RGB)] -- one baseline 3-component stream re-labelled so
    that libjpeg's colour space choice (jdapimin.c default_decompress_parms) takes each of its
    branches: JFIF marker, Adobe APP14 transform 0 / 1 / 2, component ids 1,2,3 / 'R','G','B' / other."""
    import struct
    assert data[:2] == b'\xff\xd8'
    segs = []                                   # header segments up to and including SOS
    i = 2
    while True:
        assert data[i] == 0xFF
        m = data[i + 1]
        L = struct.unpack('>H', data[i + 2:i + 4])[0]
        segs.append((m, data[i + 4:i + 2 + L]))
        i += 2 + L
        if m == 0xDA:
            break
    tail = data[i:]

    def build(keep_jfif, adobe, ids):
        out = b'\xff\xd8'
        if keep_jfif:
            body = b'JFIF\0\x01\x01\0\0\x01\0\x01\0\0'
            out += b'\xff\xe0' + struct.pack('>H', len(body) + 2) + body
        if adobe is not None:
            body = b'Adobe' + struct.pack('>HHHB', 100, 0, 0, adobe)
            out += b'\xff\xee' + struct.pack('>H', len(body) + 2) + body
        for (m, body) in segs:
            if m in (0xE0, 0xEE):
                continue
            body = bytearray(body)
            if ids is not None and m in (0xC0, 0xC1):
                for c in range(3):
                    body[6 + 3 * c] = ids[c]
            if ids is not None and m == 0xDA:
                for c in range(3):
                    body[1 + 2 * c] = ids[c]
            out += bytes([0xFF, m]) + struct.pack('>H', len(body) + 2) + bytes(body)
        return out + tail

    return [
        ('jfif', build(True, None, None), False),
        ('jfif+adobe0', build(True, 0, None), False),            # JFIF wins
        ('jfif+ids_rgb', build(True, None, (82, 71, 66)), False),
        ('adobe0', build(False, 0, None), True),
        ('adobe1', build(False, 1, None), False),
        ('adobe2', build(False, 2, None), False),                # unknown transform: warning, YCbCr
        ('adobe1+ids_rgb', build(False, 1, (82, 71, 66)), False),
        ('ids_123', build(False, None, (1, 2, 3)), False),
        ('ids_rgb', build(False, None, (82, 71, 66)), True),
        ('ids_other', build(False, None, (0, 1, 2)), False),
    ]


def write_params_file(gs, directory, image_glob='*.png'):
    """params.yml + dials.png for golden set `gs` in `directory` (no reference tree needed);
    returns the path of params.yml."""
    import cv2
    import yaml
    cv2.imwrite(os.path.join(str(directory), 'dials.png'), gs.template)
    needle_data = [dict(
        name=gs.names[i], color_range=dict(zip('hls', gs.dials[i]['color_range'])),
        dist_from_center=int(gs.dist_from_center[i]), circle_thickness=int(gs.circle_thickness[i]),
        angle_of_zero=gs.dials[i]['angle_of_zero'], center=list(gs.dials[i]['center']),
        diameter=int(gs.diameters[i]), negative_momentum=gs.dials[i]['negative_momentum'])
        for i in range(len(gs.names))]
    (x0, y0, x1, y1) = gs.meter_rect
    doc = dict(image_glob=image_glob, meter_rect=dict(top_left=[x0, y0], bottom_right=[x1, y1]),
               dials_template='dials.png', dials_template_match_threshold=int(gs.match_threshold),
               dials_template_size=[gs.tmpl_w, gs.tmpl_h], hue_shift=gs.hue_shift,
               needle_color=dict(h=125, l=80, s=130), needle_color_range=dict(h=9, l=45, s=35),
               needle_data=needle_data)
    path = os.path.join(str(directory), 'params.yml')
    with open(path, 'w') as fp:
        fp.write(yaml.safe_dump(doc))
    return path
