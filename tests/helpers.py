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
        with open(os.path.join(GOLDEN, 'set%d_expected.json' % set_no)) as fp:
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


def write_params_file(gs, directory):
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
    doc = dict(image_glob='*.png', meter_rect=dict(top_left=[x0, y0], bottom_right=[x1, y1]),
               dials_template='dials.png', dials_template_match_threshold=int(gs.match_threshold),
               dials_template_size=[gs.tmpl_w, gs.tmpl_h], hue_shift=gs.hue_shift,
               needle_color=dict(h=125, l=80, s=130), needle_color_range=dict(h=9, l=45, s=35),
               needle_data=needle_data)
    path = os.path.join(str(directory), 'params.yml')
    with open(path, 'w') as fp:
        fp.write(yaml.safe_dump(doc))
    return path
