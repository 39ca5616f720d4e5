#!/usr/bin/env python
"""Times the fused colour pass (L plane + mask, 5 B/px) for each kernel implementation.
Measurement tool: `python tools/bench_convert.py [frames]`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from meterelf_b200 import _engine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
H, W = 640, 480
dev = torch.device('cuda', 0)
g = torch.Generator(device=dev)
g.manual_seed(1)
frames = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
out = {}
for impl in ('fma', 'tab'):
    os.environ['MELF_CONVERT_IMPL'] = impl
    for (name, hls, light, mask) in (('L+mask', 0, 1, 1), ('HLS', 1, 0, 0), ('HLS+L+mask', 1, 1, 1)):
        d_h = torch.empty((n, H, W, 3), dtype=torch.uint8, device=dev) if hls else None
        d_l = torch.empty((n, H, W), dtype=torch.uint8, device=dev) if light else None
        d_m = torch.empty((n, H, W), dtype=torch.uint8, device=dev) if mask else None
        stream = torch.cuda.current_stream()

        def run():
            _engine.convert_hls_device(0, frames.data_ptr(), n, H, W, W * 3, H * W * 3, 128, True,
                                       d_h.data_ptr() if hls else 0, d_l.data_ptr() if light else 0,
                                       (116, 35, 95), (134, 125, 165), d_m.data_ptr() if mask else 0,
                                       stream.cuda_stream)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(10):
            run()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        bpp = 3 + 3 * hls + light + mask
        gbs = bpp * n * H * W / ms / 1e6
        res = tuple(t.clone() if t is not None else None for t in (d_h, d_l, d_m))
        key = name
        if key in out:
            same = all((a is None) or torch.equal(a, b) for (a, b) in zip(out[key], res))
        else:
            out[key] = res
            same = None
        print('%-4s %-11s %8.3f ms  %7.1f GB/s  (%d B/px)%s' % (
            impl, name, ms, gbs, bpp, '' if same is None else '  identical to fma: %s' % same))
