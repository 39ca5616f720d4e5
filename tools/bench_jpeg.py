#!/usr/bin/env python
"""Ingest bench: JPEG bytes -> BGR frames / readings on one B200.

    python tools/bench_jpeg.py [--streams 4096] [--reps 5] [--json out.json]

Workload: the sample JPEGs of set 1 (480x640, 4:2:0, ~35 KB each; git-ignored full set when
present, else the committed subset), repeated to `--streams` streams.  Reports

* `decode`: melf_decode_jpeg (pinned staging copy + H2D + 4 kernels) per batch, per-stage CUDA-event
  times (melf_profile_*), frames/s, and the HBM view of every stage (algorithmic bytes / time
  against MEASURED_PEAKS.json);
* `e2e`: melf_read_jpeg_batch, host JPEG bytes in -> host readings out, frames/s;
* `cpu`: the oracle's decode and (when cv2 is importable on the box) cv2.imdecode, ms per file on
  one core.

The first frames of the batch are checked against the oracle before timing.
Measurement tool, not product code.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import bench  # noqa: E402
import oracle  # noqa: E402
from helpers import GOLDEN, golden_set  # noqa: E402
from test_jpeg_oracle import load_streams  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--streams', type=int, default=4096)
    ap.add_argument('--reps', type=int, default=5)
    ap.add_argument('--json', default='')
    ap.add_argument('--skip-decode', action='store_true', help='only the e2e leg (large batches)')
    args = ap.parse_args()
    from meterelf_b200 import _engine
    full = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    streams = load_streams(full if os.path.exists(full) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
    base = [s for s in streams if s[4] == 1 and s[2] == (640, 480)]
    (h, w) = base[0][2]
    data = [base[i % len(base)][1] for i in range(args.streams)]
    n = len(data)
    (blob, offsets) = _engine.pack_streams(data)
    pinned = torch.from_numpy(blob.copy()).pin_memory()
    blob_p = pinned.numpy()
    gs = golden_set(1)
    eng = gs.engine(cropped_input=False)
    peaks = bench.load_peaks()

    rep = {'workload': '%d streams %dx%d 4:2:0, %.1f KB avg' % (n, w, h, int(offsets[-1]) / n / 1024)}
    if not args.skip_decode:
        out = torch.zeros((n, h, w, 3), dtype=torch.uint8, device='cuda:0')
        st = torch.cuda.current_stream()

        def decode():
            eng.decode_jpeg_device(blob_p, offsets, h, w, out.data_ptr(), w * 3, h * w * 3, st.cuda_stream)
        for _ in range(2):
            decode()
        torch.cuda.synchronize()
        for i in (0, 1, len(base) - 1):
            assert np.array_equal(out[i].cpu().numpy(), oracle.decode_jpeg(data[i])), i
        eng.profile_enable(True)
        t0 = time.perf_counter()
        for _ in range(args.reps):
            decode()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / args.reps
        prof = eng.profile_read()
        eng.profile_enable(False)
        stages = {k: v[0] / max(1, v[1]) for (k, v) in prof.items() if k.startswith('jpeg_')}
        blocks = (w // 8) * (h // 8) * 3 // 2
        jpeg_bytes = int(offsets[-1])
        algo = {                                             # algorithmic bytes per batch, per stage
            'jpeg_upload': jpeg_bytes + n * blocks * 128,    # H2D + coefficient clear
            'jpeg_destuff': 2 * jpeg_bytes,
            'jpeg_huffman': jpeg_bytes,                      # latency-bound: bytes are not the limit
            'jpeg_idct': n * blocks * (128 + 64),
            'jpeg_color': n * (blocks * 64 + h * w * 3),
        }
        dev_ms = sum(stages.values())
        rep.update({
               'decode': {'wall_ms_per_batch': wall * 1e3, 'device_ms_per_batch': dev_ms,
                          'frames_per_s_device': n / (dev_ms / 1e3), 'frames_per_s_wall': n / wall,
                          'stages_ms': stages,
                          'stages_gbs': {k: algo[k] / (stages[k] / 1e3) / 1e9 for k in stages if stages[k] > 0},
                          'hbm_peak_gbs': peaks['hbm_gbs'], 'peak_source': peaks['source']}})
        print(json.dumps(rep['decode'], indent=1))
        del out
        torch.cuda.empty_cache()

    # end to end: host JPEG bytes -> host readings
    res = eng.read_jpeg_streams(data[:64])               # warm-up + check
    crops = np.stack([oracle.decode_jpeg(d)[gs.meter_rect[1]:gs.meter_rect[3], gs.meter_rect[0]:gs.meter_rect[2]]
                      for d in data[:8]])
    want = oracle.read_batch(gs.oracle_params(), np.ascontiguousarray(crops), n_threads=4)
    assert all(int(a['status']) == int(b['status']) and float(a['value']) == float(b['value']) or
               (a['value'] != a['value'] and b['value'] != b['value']) for (a, b) in zip(res, want))
    lib = eng.lib
    import ctypes as C
    from meterelf_b200._engine import RESULT_DTYPE
    results = np.zeros((n,), RESULT_DTYPE)
    for _ in range(2):
        eng._check(lib.melf_read_jpeg_batch(eng._ctx, blob_p.ctypes.data, offsets.ctypes.data, n, results.ctypes.data))
    t0 = time.perf_counter()
    for _ in range(args.reps):
        eng._check(lib.melf_read_jpeg_batch(eng._ctx, blob_p.ctypes.data, offsets.ctypes.data, n, results.ctypes.data))
    e2e = (time.perf_counter() - t0) / args.reps
    jpeg_bytes = int(offsets[-1])
    eng.profile_enable(True)
    eng._check(lib.melf_read_jpeg_batch(eng._ctx, blob_p.ctypes.data, offsets.ctypes.data, n, results.ctypes.data))
    prof = eng.profile_read()
    eng.profile_enable(False)
    rep['e2e_stages_ms'] = {k: v[0] for (k, v) in prof.items() if v[1]}       # summed over the chunks of one call
    rep['e2e'] = {'ms_per_batch': e2e * 1e3, 'frames_per_s': n / e2e, 'h2d_bytes_per_batch': jpeg_bytes,
                  'd2h_bytes_per_batch': int(results.nbytes),
                  'frames_read_ok': int((results['status'] == 0).sum())}
    print(json.dumps(rep['e2e'], indent=1))

    # CPU: oracle port and, when present, the real thing
    k = 24
    t0 = time.perf_counter()
    for d in data[:k]:
        oracle.decode_jpeg(d)
    rep['cpu'] = {'oracle_ms_per_file': (time.perf_counter() - t0) / k * 1e3, 'cores': 1}
    try:
        import cv2
        cv2.setNumThreads(1)
        t0 = time.perf_counter()
        for d in data[:4 * k]:
            cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR)
        rep['cpu']['cv2_imdecode_ms_per_file'] = (time.perf_counter() - t0) / (4 * k) * 1e3
    except ImportError:
        pass
    print(json.dumps(rep['cpu'], indent=1))
    if args.json:
        with open(args.json, 'w') as fp:
            json.dump(rep, fp, indent=1)


if __name__ == '__main__':
    main()
