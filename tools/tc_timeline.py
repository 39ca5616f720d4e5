"""CTA 0's per-pass clock64 stamps of the locate kernel (MELF_TC_TIMELINE debug dump) on the bench's
headline workload.  Measurement tool, not product code.   python tools/tc_timeline.py [out.txt]"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
out = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/timeline.txt'
os.environ['MELF_TC_TIMELINE'] = out
import torch  # noqa: E402
import bench  # noqa: E402

tmp = tempfile.TemporaryDirectory()
wl = bench.Workload(1, tmp.name, 0)
n = 4096
dev = torch.device('cuda', 0)
d_frames = bench.make_frames_device(torch, wl, n, 640, 480, 1234, dev)
d_xy = torch.zeros((n, 2), dtype=torch.int32, device=dev)
d_val = torch.zeros((n,), dtype=torch.float32, device=dev)
st = torch.cuda.current_stream()
for i in range(3):
    wl.engine.match_template_device(d_frames.data_ptr(), 3, n, 640, 480, 480 * 3, 640 * 480 * 3, True,
                                    d_xy.data_ptr(), d_val.data_ptr(), 0, st.cuda_stream)
torch.cuda.synchronize()
rows = [[int(v) for v in line.split()] for line in open(out).read().splitlines() if line.strip()]
print('# pass: issuer_start slabs_in issuer_start+stage_wait accum_committed drain_done - ws_done next_slabs_done')
for r in rows:
    print(' '.join('%8d' % v for v in r))
if len(rows) >= 6:
    p = [rows[i + 1][0] - rows[i][0] for i in range(2, 6)]
    print('# pass period (clk):', p)
    print('# window sums:', [rows[i][6] - rows[i - 1][4] for i in range(3, 7)], ' staging:',
          [rows[i][7] - rows[i][6] for i in range(3, 7)], ' drain:', [rows[i][4] - max(rows[i][7], rows[i][3]) for i in range(3, 7)])
