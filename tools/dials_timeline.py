"""Debug probe: per-phase clock64 stamps of the dial kernel (frames 0..7, dial 0)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
os.environ['MELF_DIALS_TIMELINE'] = 'gpurun_out/dials_timeline.txt'
import torch
import bench
from helpers import golden_set
from meterelf_b200 import _engine
gs = golden_set(1)
crops = bench.valid_crops(gs)
n = 4096
dev = torch.device('cuda', 0)
eng = gs.engine(cropped_input=False, device=0)
d_frames = bench.make_frames_device(torch, gs, crops, n, 1234, dev)
d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream()
for i in range(3):
    eng.read_device(d_frames.data_ptr(), n, 640, 480, 480 * 3, 640 * 480 * 3, d_out.data_ptr(), st.cuda_stream)
torch.cuda.synchronize()
print('start color inrange close comps momentum ring(end)')
print(open('gpurun_out/dials_timeline.txt').read())
