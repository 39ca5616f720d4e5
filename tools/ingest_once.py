"""JPEG bytes -> readings rate of melf_read_jpeg_batch (bench.py's ingest leg alone): python tools/ingest_once.py [streams]
Environment switches of the ingest pipeline (MELF_JPEG_LANES / _SLOTS / _CHUNK / _SORT) are read once per process."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402
import bench  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
torch.cuda.set_device(0)
D = bench.Dist(torch, torch.device('cuda', 0))
tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(1, tmp.name, 0)
r = bench.ingest_leg(torch, D, wl, n, reps=4)
print({k: os.environ[k] for k in os.environ if k.startswith('MELF_JPEG')}, n, 'streams: %.0f frames/s, mismatches %s' % (r['value'], r['golden_line_mismatches_per_rank']))
