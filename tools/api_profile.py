"""cProfile of the drop-in call get_meter_values(params_file, filenames) over JPEG files on a RAM disk:
python tools/api_profile.py [files].  Where does the host time go?"""
import cProfile
import os
import pstats
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import bench  # noqa: E402
from helpers import GOLDEN  # noqa: E402
from test_jpeg_oracle import load_streams  # noqa: E402
import meterelf_b200  # noqa: E402

n_files = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(1, tmp.name, 0)
full = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
streams = load_streams(full if os.path.exists(full) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
base = [s for s in streams if s[4] == 1 and s[2] == (640, 480)]
d = tempfile.mkdtemp(prefix='melf_api_', dir='/dev/shm' if os.path.isdir('/dev/shm') else None)
try:
    paths = []
    for (name, data, _shape, _sha, _set) in base:
        paths.append(os.path.join(d, name))
        with open(paths[-1], 'wb') as fp:
            fp.write(data)
    names = [paths[i % len(paths)] for i in range(n_files)]
    list(meterelf_b200.get_meter_values(wl.params_file, names[:1024]))
    import numpy as np
    from meterelf_b200 import _api, _engine
    pinned_buffer = _engine._blob_buffer
    for (label, prefetch, pinned) in (('2 threads, pinned pool', 2, True), ('1 thread, pinned pool', 1, True),
                                      ('2 threads, pageable', 2, False), ('1 thread, pageable (before)', 1, False)):
        _api.PREFETCH_BATCHES = prefetch
        _engine._blob_buffer = pinned_buffer if pinned else (lambda nbytes: np.empty((nbytes,), np.uint8))
        for rep in range(3):
            t0 = time.perf_counter()
            list(meterelf_b200.get_meter_values(wl.params_file, names))
            dt = time.perf_counter() - t0
            print('%s, run %d: %.3f s, %.0f files/s' % (label, rep, dt, n_files / dt))
    _api.PREFETCH_BATCHES = 2
    _engine._blob_buffer = pinned_buffer
    pr = cProfile.Profile()
    pr.enable()
    list(meterelf_b200.get_meter_values(wl.params_file, names))
    pr.disable()
    pstats.Stats(pr).sort_stats('cumulative').print_stats(28)
finally:
    shutil.rmtree(d, ignore_errors=True)
