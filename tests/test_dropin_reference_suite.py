"""The reference's OWN test file (``tests/test_meterelf.py``, vendored unchanged into the git-ignored
``oracle/_ref/tests`` by ``__graft_entry__.build()``) run against the drop-in: ``meterelf`` in
``sys.modules`` is ``meterelf_b200``.

The project directory the file expects is assembled from the golden fixtures (the reference tree does
not exist on the GPU box): ``sample-images{1,2}/`` with every sample JPEG (``tests/golden/_full/
jpeg_all.npz``), ``params.yml`` + the dial template, and ``tests/sample-images{1,2}_stdout.txt`` from
``set{1,2}_expected.json``.

Expected outcome: everything passes except the two tests that compare the ``match val`` STRING of
``20180814021310-00-e02.jpg``: exact integer TM_CCOEFF gives 17495718.0 where OpenCV's float32 DFT
printed 17495704.0 (the one documented deviation).
"""
import os
import re
import shutil
import subprocess
import sys

import numpy as np
import pytest

from helpers import GOLDEN, golden_set, write_params_file
from test_jpeg_oracle import load_streams

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_TEST = os.path.join(ROOT, 'oracle', '_ref', 'tests', 'test_meterelf.py')
FULL_JPEGS = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')

CONFTEST = '''
import sys
sys.path.insert(0, %r)
import meterelf_b200
from meterelf_b200 import _api, _calibration, _debug, _main, _params, exceptions  # noqa
sys.modules['meterelf'] = meterelf_b200
for name in ('_api', '_calibration', '_debug', '_main', '_params', 'exceptions'):
    sys.modules['meterelf.' + name] = getattr(meterelf_b200, name)
'''

KNOWN_DEVIATIONS = {
    'test_main_with_all_sample_images[sample-images1]',        # one line: the match val string below
    'test_raises_on_debug_mode[20180814021310-00-e02.jpg]',    # 17495718.0 (exact) vs 17495704.0 (cv2 DFT)
}


@pytest.mark.skipif(not (os.path.exists(REF_TEST) and os.path.exists(FULL_JPEGS)),
                    reason='oracle/_ref/tests or the full JPEG set is absent (made by __graft_entry__.build() '
                           'where /root/reference exists)')
def test_reference_test_file_against_the_drop_in(tmp_path):
    proj = tmp_path / 'proj'
    (proj / 'tests').mkdir(parents=True)
    shutil.copy(REF_TEST, proj / 'tests' / 'test_meterelf.py')
    (proj / 'tests' / 'conftest.py').write_text(CONFTEST % ROOT)
    streams = load_streams(FULL_JPEGS)
    for set_no in (1, 2):
        gs = golden_set(set_no)
        d = proj / ('sample-images%d' % set_no)
        d.mkdir()
        write_params_file(gs, d, image_glob='*.jpg')
        lines = []
        for (name, data, _shape, _sha, s_no) in streams:
            if s_no == set_no:
                (d / name).write_bytes(data)
                lines.append('%s: %s' % (name, gs.expected[name]['line']))
        (proj / 'tests' / ('sample-images%d_stdout.txt' % set_no)).write_text('\n'.join(sorted(lines)) + '\n')
    env = dict(os.environ, PYTHONPATH=ROOT)
    env.pop('DEBUG', None)
    res = subprocess.run([sys.executable, '-m', 'pytest', '-q', '-p', 'no:cacheprovider', '-rf',
                          'tests/test_meterelf.py'], cwd=str(proj), env=env, stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True, timeout=900)
    out = res.stdout
    failed = set(re.findall(r'^FAILED tests/test_meterelf\.py::(\S+)', out, re.M))
    m = re.search(r'(\d+) passed', out)
    passed = int(m.group(1)) if m else 0
    assert failed == KNOWN_DEVIATIONS, out[-4000:]
    assert passed == 7 - len(KNOWN_DEVIATIONS), out[-2000:]          # the file holds 7 test cases
    # the one differing line of sample-images1 is the documented one, nothing else
    diff = [ln for ln in out.splitlines() if re.search(r'\.jpg\s*: got: ', ln)]
    assert any('17495718.0' in ln and '17495704.0' in ln for ln in diff), out[-4000:]
    assert all('20180814021310-00-e02.jpg' in ln for ln in diff), diff
