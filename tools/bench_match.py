#!/usr/bin/env python
"""Template-correlation micro-bench (BASELINE.json configs[4]): the sample 188x119 dial template
against single-channel search windows of growing size -- cv2.matchTemplate(TM_CCOEFF) +
cv2.minMaxLoc semantics through ``melf_match_template``.

    python tools/bench_match.py [--json out.json]

Per window size: ms per batch, frames/s, result positions/s, algorithmic TOP/s
(2 x positions x 188 x 119 per frame) and its fraction of the int8 tensor peak
(2 x the measured bf16 figure of MEASURED_PEAKS.json).  The first frame of every size is checked
against the CPU oracle's arg-max (exact integer arithmetic) before timing.
Measurement tool, not product code.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import bench  # noqa: E402
from helpers import golden_set  # noqa: E402

SIZES = [(220, 135), (250, 250), (320, 240), (640, 480), (1280, 720)]     # (w, h), SURVEY 8d


def exact_argmax(img, tmpl):
    """Exact TM_CCOEFF arg-max of one small frame with numpy int64 (first in raster order)."""
    (th, tw) = tmpl.shape
    (h, w) = img.shape
    n_t = th * tw
    sum_t = int(tmpl.astype(np.int64).sum())
    win = np.lib.stride_tricks.sliding_window_view(img.astype(np.int64), (th, tw))
    stl = np.einsum('yxij,ij->yx', win, tmpl.astype(np.int64))
    sw = win.sum(axis=(2, 3))
    num = n_t * stl - sum_t * sw
    score = (num.astype(np.float64) / n_t).astype(np.float32)
    idx = int(np.argmax(score))
    return (idx % score.shape[1], idx // score.shape[1], float(score.flat[idx]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--json', default='')
    ap.add_argument('--budget-mb', type=int, default=2048, help='bytes of frames per batch')
    args = ap.parse_args()
    peaks = bench.load_peaks()
    int8_peak = 2.0 * peaks['bf16_tflops']
    gs = golden_set(1)
    eng = gs.engine(cropped_input=True, device=0)
    tmpl = np.ascontiguousarray(gs.template)
    (th, tw) = tmpl.shape
    dev = torch.device('cuda', 0)
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    rows = []
    for (w, h) in SIZES:
        n = max(8, min(4096, args.budget_mb * (1 << 20) // (w * h)))
        frames = torch.randint(0, 256, (n, h, w), dtype=torch.uint8, device=dev, generator=g)
        # plant the template into frame 0 so that its maximum is far above the noise
        frames[0, 3:3 + th, 5:5 + tw] = torch.from_numpy(tmpl).to(dev)
        d_xy = torch.zeros((n, 2), dtype=torch.int32, device=dev)
        d_val = torch.zeros((n,), dtype=torch.float32, device=dev)
        st = torch.cuda.current_stream()

        def run():
            eng.match_template_device(frames.data_ptr(), 1, n, h, w, w, h * w, False,
                                      d_xy.data_ptr(), d_val.data_ptr(), 0, st.cuda_stream)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        got = (int(d_xy[0, 0]), int(d_xy[0, 1]), float(d_val[0]))
        chk = None
        if w * h <= 320 * 240:
            k = 1 if n > 1 else 0
            want = exact_argmax(frames[k].cpu().numpy(), tmpl)
            gk = (int(d_xy[k, 0]), int(d_xy[k, 1]), float(d_val[k]))
            chk = gk == want
            assert chk, (gk, want)
        assert got[:2] == (5, 3), got
        (e0, e1) = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        reps = 10
        e0.record(st)
        for _ in range(reps):
            run()
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        pos = (w - tw + 1) * (h - th + 1)
        tops = 2.0 * pos * tw * th * n / (ms / 1e3) / 1e12
        rows.append({'window': '%dx%d' % (w, h), 'result_map': '%dx%d' % (w - tw + 1, h - th + 1),
                     'frames': n, 'ms_per_batch': ms, 'frames_per_s': n / (ms / 1e3),
                     'positions_per_s': pos * n / (ms / 1e3), 'algorithmic_tops': tops,
                     'frac_of_int8_peak': tops / int8_peak, 'exact_argmax_check': chk})
        print('%-9s map %-9s n=%-5d %8.3f ms  %10.0f frames/s  %7.1f TOP/s  %.3f of int8 peak  check=%s'
              % (rows[-1]['window'], rows[-1]['result_map'], n, ms, n / (ms / 1e3), tops,
                 tops / int8_peak, chk))
        del frames
    out = {'template': '%dx%d' % (tw, th), 'int8_peak_tops': int8_peak, 'peak_source': peaks['source'],
           'rows': rows}
    if args.json:
        with open(args.json, 'w') as fp:
            json.dump(out, fp, indent=1)


if __name__ == '__main__':
    main()
