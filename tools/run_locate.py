"""The locate kernel alone on the bench's headline frames: python tools/run_locate.py [launches].
Measurement helper (ncu target, timeline runs); prints the CUDA-event time per launch."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402
import bench  # noqa: E402

launches = int(sys.argv[1]) if len(sys.argv) > 1 else 5
tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(int(os.environ.get('MELF_SET', '1')), tmp.name, 0)
n = 4096
dev = torch.device('cuda', 0)
d_frames = bench.make_frames_device(torch, wl, n, 640, 480, 1234, dev)
d_xy = torch.zeros((n, 2), dtype=torch.int32, device=dev)
d_val = torch.zeros((n,), dtype=torch.float32, device=dev)
st = torch.cuda.current_stream()
(e0, e1) = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
for i in range(launches):
    if i == launches - 1:
        e0.record(st)
    wl.engine.match_template_device(d_frames.data_ptr(), 3, n, 640, 480, 480 * 3, 640 * 480 * 3, True,
                                    d_xy.data_ptr(), d_val.data_ptr(), 0, st.cuda_stream)
e1.record(st)
torch.cuda.synchronize()
print('locate + unpack, last launch: %.4f ms' % e0.elapsed_time(e1))
