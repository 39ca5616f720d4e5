"""Per-stage GPU time of melf_read_jpeg_batch (CUDA events around every launch; the launches of different chunks
overlap, so the sum exceeds the wall time): python tools/ingest_stages.py [streams]"""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from helpers import GOLDEN  # noqa: E402
from test_jpeg_oracle import load_streams  # noqa: E402
from meterelf_b200 import _engine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(1, tmp.name, 0)
eng = wl.engine
full = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
streams = load_streams(full if os.path.exists(full) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
base = [s for s in streams if s[4] == 1 and s[2] == (640, 480)]
(blob, offsets) = _engine.pack_streams([base[i % len(base)][1] for i in range(n)])
blob_p = torch.from_numpy(blob.copy()).pin_memory().numpy()
results = np.zeros((n,), _engine.RESULT_DTYPE)


def call():
    eng._check(eng.lib.melf_read_jpeg_batch(eng._ctx, blob_p.ctypes.data, offsets.ctypes.data, n, results.ctypes.data))


for _ in range(2):
    call()
t0 = time.perf_counter()
call()
wall = time.perf_counter() - t0
eng.profile_enable(True)
call()
prof = eng.profile_read()
eng.profile_enable(False)
print('%d streams: %.1f ms wall (%.0f frames/s); summed GPU ms per stage:' % (n, wall * 1e3, n / wall),
      {k: round(v[0], 2) for (k, v) in prof.items() if v[1]}, 'launches', {k: v[1] for (k, v) in prof.items() if v[1]})
