"""Debug probe: per-phase clock64 stamps of the dial kernel (frames 0..7, dial 0) on the bench's headline
workload.  Measurement tool, not product code.   python tools/dials_timeline.py"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
os.environ['MELF_DIALS_TIMELINE'] = 'gpurun_out/dials_timeline.raw'
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from meterelf_b200 import _engine  # noqa: E402

tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(int(os.environ.get('MELF_SET', '1')), tmp.name, 0)
n = 4096
dev = torch.device('cuda', 0)
d_frames = bench.make_frames_device(torch, wl, n, 640, 480, 1234, dev)
d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream()
for i in range(3):
    wl.engine.read_device(d_frames.data_ptr(), n, 640, 480, 480 * 3, 640 * 480 * 3, d_out.data_ptr(), st.cuda_stream)
torch.cuda.synchronize()
rows = np.array([[int(v) for v in line.split()] for line in open('gpurun_out/dials_timeline.raw').read().splitlines() if line.strip()])
print('# dial 0 of frames 0..7: clocks after the start: dial colour, inRange, closing, components, momentum, ring + mean (end); [7] = window staged')
print(rows)
d = np.diff(rows[:, :7], axis=1)
names = ['staging + colour', 'inRange', 'closing', 'components', 'momentum', 'ring + mean']
mean = d.mean(axis=0)
print('# mean clocks per phase:', {k: int(v) for (k, v) in zip(names, mean)}, 'total', int(mean.sum()))
print('# share:', {k: round(float(v / mean.sum()), 3) for (k, v) in zip(names, mean)})
