import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
os.environ['MELF_TC_TIMELINE'] = 'gpurun_out/timeline.txt'
import torch, numpy as np
import bench
from helpers import golden_set
from meterelf_b200 import _engine
gs = golden_set(1)
crops = bench.valid_crops(gs)
n = 4096
dev = torch.device('cuda', 0)
eng = gs.engine(cropped_input=False, device=0)
d_frames = bench.make_frames_device(torch, gs, crops, n, 1234, dev)
d_xy = torch.zeros((n, 2), dtype=torch.int32, device=dev); d_val = torch.zeros((n,), dtype=torch.float32, device=dev)
st = torch.cuda.current_stream()
for i in range(3):
    eng.match_template_device(d_frames.data_ptr(), 3, n, 640, 480, 480*3, 640*480*3, True, d_xy.data_ptr(), d_val.data_ptr(), 0, st.cuda_stream)
torch.cuda.synchronize()
print(open('gpurun_out/timeline.txt').read())
