"""The whole path (locate + dials) on the bench's headline frames: python tools/run_path.py [steps].
Measurement helper (ncu target); prints the CUDA-event time per step."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402
import bench  # noqa: E402
from meterelf_b200 import _engine  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(int(os.environ.get('MELF_SET', '1')), tmp.name, 0)
n = 4096
dev = torch.device('cuda', 0)
d_frames = bench.make_frames_device(torch, wl, n, 640, 480, 1234, dev)
d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream()
(e0, e1) = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
for i in range(steps):
    if i == steps - 1:
        e0.record(st)
    wl.engine.read_device(d_frames.data_ptr(), n, 640, 480, 480 * 3, 640 * 480 * 3, d_out.data_ptr(), st.cuda_stream)
e1.record(st)
torch.cuda.synchronize()
res = d_out.cpu().numpy().view(_engine.RESULT_DTYPE).reshape(n)
print('whole path, last step: %.4f ms, frames read ok: %d' % (e0.elapsed_time(e1), int((res['status'] == 0).sum())))
