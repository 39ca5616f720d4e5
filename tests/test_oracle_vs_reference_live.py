"""The oracle port against the UNMODIFIED reference, run live on inputs no fixture holds.

``oracle/_ref`` is the reference package as it lies in ``/root/reference`` (copied by
``__graft_entry__.build()``); ``oracle.RefPool`` runs it in worker processes under the three environment shims
of SURVEY.md 8c.  Here: the committed sample meter crops, perturbed with fresh random gain / offset / noise per
seed (so every frame still shows a meter and the template match is unambiguous, but no pixel equals a fixture's),
read by the reference and by the port -- status, failing / unreadable dials and the float64 dial positions and
value must agree bit for bit.  CPU only; skipped where ``oracle/_ref`` has not been built.
"""
import numpy as np
import pytest

import oracle
from helpers import golden_set, write_params_file


def perturbed_crops(gs, seed, copies):
    crops = gs.subset()['crops']
    rng = np.random.default_rng(seed)
    out = []
    for c in range(copies):
        for crop in crops:
            gain = rng.uniform(0.85, 1.15)
            offset = rng.uniform(-12, 12)
            noise = rng.normal(0.0, rng.uniform(0.0, 5.0), crop.shape)
            shift = rng.integers(-3, 4, size=2)
            img = np.clip(crop.astype(np.float64) * gain + offset + noise, 0, 255).astype(np.uint8)
            img = np.roll(img, (int(shift[0]), int(shift[1])), axis=(0, 1))
            out.append(img)
    return np.ascontiguousarray(np.stack(out))


@pytest.mark.parametrize('set_no,seed', [(1, 101), (2, 202)])
def test_port_equals_the_reference_on_perturbed_sample_crops(set_no, seed, tmp_path):
    if not oracle.ref_available():
        pytest.skip('oracle/_ref not built (needs /root/reference at build time)')
    gs = golden_set(set_no)
    params_file = write_params_file(gs, tmp_path)
    crops = perturbed_crops(gs, seed, copies=2)
    with oracle.RefPool(params_file, crops, workers=4) as pool:
        (_wall, ref_records) = pool.run()
    assert len(ref_records) == len(crops)
    O = gs.oracle_params()
    got = np.zeros((len(crops),), oracle.RESULT_DTYPE)
    for (i, crop) in enumerate(crops):
        (res, _taps) = oracle.read_frame(O, crop)
        got[i] = res
    statuses = sorted(set(int(r['status']) for r in ref_records))
    assert 0 in statuses                                   # most frames still read
    bad = oracle.ref_mismatches(ref_records, got, len(gs.names))
    assert bad == 0, (bad, len(crops), statuses)
