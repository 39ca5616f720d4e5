#!/usr/bin/env python
"""Throughput of the dial-reading hot path: meter readings/sec on 640x480 frames.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the whole path (locate dials + read 4 dials + value) over one
HBM-resident synthetic batch (default 4096 frames of 640x480x3 u8 per GPU = 3.77 GB,
larger than L2, so no flush between steps; --size 1280x720 gives BASELINE.json's second
throughput config, 32768 frames split over the GPUs).  Frames are "S-planted" (SURVEY.md 8d):
uniform noise canvases with decoded sample meter crops pasted at the meter_rect with a
seeded +-6 px jitter, so every stage runs on every frame and the expected readings are
known.  Prints ONE JSON line (rank 0).  Frames shard by index over the GPUs: no data-path
collective; torch.distributed is used only for the barrier and the max-over-ranks time.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

METRIC = 'meter readings/sec (640x480 frames)'
UNIT = 'frames/s'
FRAME_H, FRAME_W = 640, 480        # default: the sample frames are portrait, 480 wide x 640 tall
SMI_QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')


def load_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as fp:
            d = json.load(fp)
        return dict(hbm_gbs=d['hbm_gbs'], bf16_tflops=d['bf16_tflops'], source='measured')
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, source='fallback')


def set_frame_size(size):
    global FRAME_H, FRAME_W
    (w, h) = (int(v) for v in size.lower().split('x'))
    (FRAME_H, FRAME_W) = (h, w)


def workload_config(frames):
    return {
        'workload': 'synthetic %dx%d (w x h) BGR batch of %d frames per GPU, S-planted sample meter '
                    'crops (sample-images1 params, meter_rect 250x250, template 188x119)'
                    % (FRAME_W, FRAME_H, frames),
        'frames_per_gpu': frames, 'frame_h': FRAME_H, 'frame_w': FRAME_W,
        'bytes_per_frame': FRAME_H * FRAME_W * 3,
        'l2': 'batch (%.2f GB) is larger than L2; no flush between steps' % (
            frames * FRAME_H * FRAME_W * 3 / 1e9),
        'parallelism': 'frames sharded by index, one process per GPU, host-side gather',
    }


def planted_offsets(n, seed):
    rng = np.random.default_rng(seed)
    return rng.integers(-6, 7, size=(n, 2))


def make_frames_host(gs, crops, n, seed):
    """Host (numpy) version of the synthetic batch, used for the CPU arms and checks."""
    rng = np.random.default_rng(seed + 1)
    (x0, y0, x1, y1) = gs.meter_rect
    frames = rng.integers(0, 256, size=(n, FRAME_H, FRAME_W, 3), dtype=np.uint8)
    off = planted_offsets(n, seed)
    for i in range(n):
        (dx, dy) = off[i]
        frames[i, y0 + dy:y1 + dy, x0 + dx:x1 + dx] = crops[i % len(crops)]
    return frames


def make_frames_device(torch, gs, crops, n, seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed + 1)
    frames = torch.randint(0, 256, (n, FRAME_H, FRAME_W, 3), dtype=torch.uint8, device=device,
                           generator=g)
    d_crops = torch.from_numpy(np.ascontiguousarray(crops)).to(device)
    (x0, y0, x1, y1) = gs.meter_rect
    off = planted_offsets(n, seed)
    for i in range(n):
        (dx, dy) = (int(off[i][0]), int(off[i][1]))
        frames[i, y0 + dy:y1 + dy, x0 + dx:x1 + dx] = d_crops[i % len(crops)]
    return frames


def valid_crops(gs):
    sub = gs.subset()
    names = [str(s) for s in sub['names']]
    keep = [i for (i, fn) in enumerate(names) if gs.expected[fn]['status'] == 0]
    return sub['crops'][keep]


class NvmlSampler:
    """SM clock, power and throttle reasons polled in-process through NVML every ~2 ms, so that a
    timed region of a few tens of milliseconds still gets a useful number of samples (nvidia-smi
    cannot loop faster than a few samples per second)."""

    REASONS = (('hw_slowdown', 0x8), ('hw_thermal_slowdown', 0x40), ('sw_thermal_slowdown', 0x20),
               ('sw_power_cap', 0x4))

    def __init__(self, gpu_index):
        import threading
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        self.rows = []
        self.done = threading.Event()
        self.thread = threading.Thread(target=self._poll, daemon=True)
        self.thread.start()

    def _poll(self):
        nv = self.nv
        while not self.done.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                power = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((sm, power, mask))
            except Exception:
                pass
            self.done.wait(0.002)

    def stop(self):
        self.done.set()
        self.thread.join(timeout=2)
        out = {'sm_mhz': None, 'sm_max_mhz': self.smax, 'reasons': [], 'samples': len(self.rows),
               'sampler': 'nvml, 2 ms period'}
        if self.rows:
            pmax = max(r[1] for r in self.rows)
            busy = [r[0] for r in self.rows if r[1] >= 0.5 * pmax] or [r[0] for r in self.rows]
            mask = 0
            for r in self.rows:
                mask |= r[2]
            out.update(sm_mhz=float(np.median(busy)), power_w_max=pmax,
                       reasons=sorted(name for (name, bit) in self.REASONS if mask & bit))
        return out


def start_clock_sampler(gpu_index):
    try:
        return NvmlSampler(gpu_index)
    except Exception:
        return ClockSampler(gpu_index)


class ClockSampler:
    def __init__(self, gpu_index):
        self.tmp = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(gpu_index), '--query-gpu=' + SMI_QUERY,
                 '--format=csv,noheader,nounits', '-lms', '100'],
                stdout=self.tmp, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        sm, smax, power, reasons = [], [], [], set()
        for line in self.tmp.read().splitlines():
            cols = [c.strip() for c in line.split(',')]
            if len(cols) < 9:
                continue
            try:
                sm.append(float(cols[1])); smax.append(float(cols[2])); power.append(float(cols[3]))
            except ValueError:
                continue
            for (name, val) in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown',
                                    'sw_power_cap'), cols[5:9]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        os.unlink(self.tmp.name)
        if sm:
            busy = [s for (s, p) in zip(sm, power) if p >= 0.5 * max(power)] or sm
            out.update(sm_mhz=float(np.median(busy)), sm_max_mhz=max(smax), samples=len(sm),
                       power_w_max=max(power), reasons=sorted(reasons))
        return out


def cpu_arm(gs, frames_host, n_threads):
    """The oracle port on host cores over full frames (meter_rect window by pointer math)."""
    import oracle
    (x0, y0, x1, y1) = gs.meter_rect
    view = frames_host[:, y0:y1, x0:x1]
    O = gs.oracle_params()
    t0 = time.perf_counter()
    res = oracle.read_batch(O, view, n_threads=n_threads)
    dt = time.perf_counter() - t0
    return (res, dt)


_RESULT_FD = None


def protect_stdout():
    """stdout carries exactly one JSON line: keep the real stdout aside and point fd 1 at stderr,
    so that nothing a library prints there (NCCL's version banner, C-level printf) can precede it."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    text = (json.dumps(line) + '\n').encode()
    if _RESULT_FD is None:
        sys.stdout.write(text.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, text)


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference itself is
    Python + OpenCV and cannot travel to the GPU box) on all host cores, same workload."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from helpers import golden_set
    gs = golden_set(1)
    cores = len(os.sched_getaffinity(0))
    sample = max(cores * 64, 256)              # ~1.5 s of CPU work per step
    frames = make_frames_host(gs, valid_crops(gs), sample, seed=1234)
    times = []
    for i in range(args.warmup + args.steps):
        (_res, dt) = cpu_arm(gs, frames, cores)
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = sample * len(times) / total
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT,
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * total / len(times), 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'u8/int32/f64', 'data': 'synthetic',
        'config': workload_config(args.frames),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': '%d frames per step' % sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    emit(line)


def load_traffic():
    """dram__bytes_read + dram__bytes_write per launch from the committed ncu capture
    (tools/ncu_summary.py traffic ... > profiles/r01_traffic.json); None when absent."""
    path = os.path.join(ROOT, 'profiles', 'r01_traffic.json')
    if os.path.exists(path):
        with open(path) as fp:
            return json.load(fp)
    return {}


def ingest_leg(torch, eng, gs, n_streams, reps=3):
    """SURVEY 8f-1, reported beside the headline: the reference's real input is JPEG files, so this
    leg times melf_read_jpeg_batch -- host JPEG bytes in (page-locked), host readings out, decode
    bit-identical to cv2.imread on the GPU -- over the portrait sample files of set 1 repeated to
    n_streams, and checks every reading against the reference's golden line."""
    from helpers import GOLDEN, format_result_line
    from test_jpeg_oracle import load_streams
    from meterelf_b200 import _engine
    full = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    streams = load_streams(full if os.path.exists(full) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
    base = [s for s in streams if s[4] == 1 and s[2] == (640, 480)]
    data = [base[i % len(base)][1] for i in range(n_streams)]
    (blob, offsets) = _engine.pack_streams(data)
    blob_p = torch.from_numpy(blob.copy()).pin_memory().numpy()
    results = np.zeros((n_streams,), _engine.RESULT_DTYPE)

    def call():
        eng._check(eng.lib.melf_read_jpeg_batch(eng._ctx, blob_p.ctypes.data, offsets.ctypes.data,
                                                n_streams, results.ctypes.data))
    for _ in range(2):
        call()
    launches0 = _engine.launch_count()
    t0 = time.perf_counter()
    for _ in range(reps):
        call()
    dt = (time.perf_counter() - t0) / reps
    launches = (_engine.launch_count() - launches0) // reps
    bad = 0
    for i in range(len(base)):
        exp = gs.expected[base[i][0]]
        got = format_result_line(results[i], gs.names)
        bad += (got != exp['line']) if exp['status'] == 0 else (got.split(' (match val')[0] != exp['line'].split(' (match val')[0])
    return {'metric': 'meter readings/sec from JPEG bytes (480x640 4:2:0 files, %.1f KB avg)' % (int(offsets[-1]) / n_streams / 1024),
            'value': n_streams / dt, 'unit': UNIT, 'streams_per_step': n_streams, 'steps': reps,
            'ms_per_step': dt * 1e3, 'h2d_bytes_per_step': int(offsets[-1]),
            'd2h_bytes_per_step': int(results.nbytes), 'gpu_launches_per_step': int(launches),
            'golden_lines_checked': len(base), 'golden_line_mismatches': int(bad),
            'note': 'melf_read_jpeg_batch: header walk on the host, Huffman / IDCT / upsampling / colour on the '
                    'GPU (bit-identical to cv2.imread), then the same locate and dial kernels; GPU-bound '
                    '(Huffman stage), see DESIGN.md section 8'}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from helpers import golden_set
    from meterelf_b200 import _engine

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise _engine.EngineUnavailableError('bench needs a CUDA device; there is no CPU fallback')
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    if world > 1:
        # NCCL's version banner goes to stdout by default; stdout carries exactly one JSON line
        os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
        dist.init_process_group('nccl', device_id=device)

    gs = golden_set(1)
    crops = valid_crops(gs)
    n = args.frames
    seed = 1234 + rank
    eng = gs.engine(cropped_input=False, device=local)
    d_frames = make_frames_device(torch, gs, crops, n, seed, device)
    d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device=device)
    stream = torch.cuda.current_stream()
    row_stride, frame_stride = FRAME_W * 3, FRAME_H * FRAME_W * 3

    def step():
        eng.read_device(d_frames.data_ptr(), n, FRAME_H, FRAME_W, row_stride, frame_stride,
                        d_out.data_ptr(), stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        step()
    barrier()
    sampler = start_clock_sampler(local) if rank == 0 else None
    launches0 = _engine.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    launches = _engine.launch_count() - launches0
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * n * args.steps / (ms_total / 1e3)

    # ---- correctness spot check of the timed output (not timed) ------------------------
    res = d_out.cpu().numpy().view(_engine.RESULT_DTYPE).reshape(n)
    n_ok = int(np.sum(res['status'] == 0))

    # ---- per-kernel durations: the same steps again with CUDA events around every launch --
    ksteps = max(3, min(args.steps, 20))
    eng.profile_enable(True)
    for _ in range(ksteps):
        step()
    prof = eng.profile_read()
    eng.profile_enable(False)
    kernel_ms = {k: (v[0] / ksteps) for (k, v) in prof.items() if v[1]}
    kernel_launches = {k: v[1] // ksteps for (k, v) in prof.items() if v[1]}

    # ---- stage 1 as a standalone fused pass over the full frames (5 B/px) ---------------
    nconv = min(n, 1024)
    d_l = torch.empty((nconv, FRAME_H, FRAME_W), dtype=torch.uint8, device=device)
    d_m = torch.empty((nconv, FRAME_H, FRAME_W), dtype=torch.uint8, device=device)

    def convert_only():
        _engine.convert_hls_device(local, d_frames.data_ptr(), nconv, FRAME_H, FRAME_W, row_stride,
                                   frame_stride, gs.hue_shift, True, 0, d_l.data_ptr(),
                                   (116, 35, 95), (134, 125, 165), d_m.data_ptr(), stream.cuda_stream)
    for _ in range(3):
        convert_only()
    torch.cuda.synchronize()
    csteps = max(5, min(args.steps, 50))
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record(stream)
    for _ in range(csteps):
        convert_only()
    c1.record(stream)
    torch.cuda.synchronize()
    conv_ms = c0.elapsed_time(c1) / csteps
    clocks = sampler.stop() if sampler else None
    del d_l, d_m

    # ---- end to end through the public host API: pinned host frames in, results out -----
    n_e2e = min(n, args.e2e_frames)
    h_frames = torch.empty((n_e2e, FRAME_H, FRAME_W, 3), dtype=torch.uint8).pin_memory()
    h_frames.copy_(d_frames[:n_e2e])
    h_np = h_frames.numpy()
    eng.read_frames(h_np[:64])
    esteps = max(2, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter()
    for _ in range(esteps):
        h_res = eng.read_frames(h_np)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    te = torch.tensor([dt], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    (x0, y0, x1, y1) = gs.meter_rect
    e2e = {'value': world * n_e2e * esteps / float(te.item()), 'unit': UNIT,
           'h2d_bytes_per_step': int(n_e2e * (y1 - y0) * (x1 - x0) * 3),
           'd2h_bytes_per_step': int(n_e2e * _engine.RESULT_DTYPE.itemsize),
           'frames_per_step': n_e2e, 'steps': esteps,
           'note': 'melf_read_batch: pinned host frames, only meter_rect rows cross PCIe '
                   '(PCIe-bound: profiles/r01_pcie_probe.txt measures 50.4 GB/s for this copy shape)'}
    assert np.array_equal(h_res['status'], res['status'][:n_e2e])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- CPU baseline (oracle port) on a bounded sample of the same frames -------------
    cores = len(os.sched_getaffinity(0))
    sample = min(n, args.cpu_frames)
    host_sample = d_frames[:sample].cpu().numpy()
    (cpu_res, cpu_dt) = cpu_arm(gs, host_sample, cores)
    mism = 0
    for i in range(sample):
        a, b = res[i], cpu_res[i]
        same = (a['status'] == b['status'] and a['match_x'] == b['match_x'] and
                a['match_y'] == b['match_y'] and
                np.array_equal(a['dial_pos'][:4], b['dial_pos'][:4], equal_nan=True))
        mism += 0 if same else 1

    peaks = load_peaks()
    traffic = load_traffic()
    step_ms = ms_total / args.steps
    rw, rh = 250 - 188 + 1, 250 - 119 + 1
    macs = rw * rh * 188 * 119
    tc_ms = kernel_ms.get('match_tc')
    int8_peak = 2.0 * peaks['bf16_tflops']
    fused = not kernel_ms.get('match_prep')
    overlap = os.environ.get('MELF_MATCH_OVERLAP', '1') not in ('0', '')
    tc_name = ('melf_match_fused_ov_kernel' if overlap else 'melf_match_fused_kernel') if fused else 'melf_match_tc_kernel'
    if tc_ms:
        tops = 2.0 * macs * n / (tc_ms / 1e3) / 1e12
        crop_bytes = 250 * 250 * 3 * n
        roofline = {
            'kernel': tc_name + (' (BGR crop -> lightness + window sums in the shadow of the MMAs, drain under the next pass; ' if fused else ' (') +
                      'TM_CCOEFF on tcgen05 kind::i8, exact integer correlation + arg-max)',
            'bound': 'tensor', 'achieved': tops, 'peak': int8_peak, 'unit': 'TOP/s',
            'frac': tops / int8_peak, 'traffic': traffic.get(tc_name),
            'ms_per_launch': tc_ms,
            'hbm_view': {'algorithmic_bytes': crop_bytes, 'achieved_gbs': crop_bytes / (tc_ms / 1e3) / 1e9,
                         'frac_of_hbm_peak': crop_bytes / (tc_ms / 1e3) / 1e9 / peaks['hbm_gbs'],
                         'note': 'the only DRAM traffic of the fused kernel is one read of the meter_rect crop '
                                 '(187500 B/frame); it is far from HBM-bound'} if fused else None,
            'note': 'algorithmic 2*%d op/frame x %d frames per launch (SURVEY 8d: 63x132 positions x 22372 MAC); '
                    'peak = 2 x %s bf16 dense (int8:bf16 tensor ratio); issued MACs are 1.50x the algorithmic '
                    'ones (K 188->256, N 133->144, M 63->64); kernel share of step %.2f'
                    % (macs, n, peaks['source'], tc_ms / step_ms),
        }
    else:
        roofline = None
    conv_bytes = 5.0 * FRAME_H * FRAME_W * nconv
    conv_gbs = conv_bytes / (conv_ms / 1e3) / 1e9
    stage = [{
        'kernel': 'melf_convert_tab_kernel (fused BGR->HLS + hue shift + L plane + mask, full frames)',
        'bound': 'hbm', 'achieved': conv_gbs, 'peak': peaks['hbm_gbs'], 'unit': 'GB/s',
        'frac': conv_gbs / peaks['hbm_gbs'], 'traffic': traffic.get('melf_convert_tab_kernel'),
        'ms_per_launch': conv_ms,
        'note': 'algorithmic 5 B/px x %d px; peak %s' % (FRAME_H * FRAME_W * nconv, peaks['source'])}]
    if kernel_ms.get('match_prep'):
        prep_bytes = n * (250 * 250 * 3 + 67584 + 64 * 144 * 4)
        gbs = prep_bytes / (kernel_ms['match_prep'] / 1e3) / 1e9
        stage.append({
            'kernel': 'melf_match_prep_kernel (meter_rect BGR -> lightness slab + window sums)',
            'bound': 'hbm', 'achieved': gbs, 'peak': peaks['hbm_gbs'], 'unit': 'GB/s',
            'frac': gbs / peaks['hbm_gbs'], 'traffic': traffic.get('melf_match_prep_kernel'),
            'ms_per_launch': kernel_ms['match_prep'],
            'note': 'algorithmic 187500 B in + 67584 + 36864 B out per frame; share of step %.2f'
                    % (kernel_ms['match_prep'] / step_ms)})
    if kernel_ms.get('dials'):
        stage.append({
            'kernel': 'melf_dials_kernel (per-dial masks, contour selection, angles, value)',
            'bound': 'issue/latency (SURVEY 8d stage 3)', 'ms_per_launch': kernel_ms['dials'],
            'frames_per_s': n / (kernel_ms['dials'] / 1e3),
            'traffic': traffic.get('melf_dials_kernel'),
            'note': 'share of step %.2f' % (kernel_ms['dials'] / step_ms)})
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
        'warmup': warmup, 'ms_per_step': step_ms,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'u8/int32 (correlation), f32 (HLS), f64 (angles)', 'data': 'synthetic',
        'config': workload_config(n),
        'e2e': e2e, 'gpu_launches': int(launches), 'clocks': clocks,
        'roofline': roofline, 'stage_rooflines': stage,
        'kernel_ms_per_step': kernel_ms, 'kernel_launches_per_step': kernel_launches,
        'cpu_baseline': {'value': sample / cpu_dt, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': '%d frames of the same batch' % sample,
                         'mismatches_vs_gpu': mism},
        'check': {'frames_read_ok': n_ok, 'frames': n},
    }
    if world == 1 and args.ingest_streams > 0 and (FRAME_H, FRAME_W) == (640, 480):
        line['ingest'] = ingest_leg(torch, eng, gs, args.ingest_streams)
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--size', default='480x640', help='frame size WxH (sample frames are 480x640; '
                    'BASELINE.json also names 1280x720)')
    ap.add_argument('--frames', type=int, default=0, help='frames per GPU (default 4096; '
                    '32768 / n_gpus for 1280x720)')
    ap.add_argument('--e2e-frames', type=int, default=4096)
    ap.add_argument('--cpu-frames', type=int, default=8192, help='CPU baseline sample (frames)')
    ap.add_argument('--ingest-streams', type=int, default=32768,
                    help='JPEG streams per step of the ingest leg (N=1 only; 0 = skip)')
    args = ap.parse_args()
    protect_stdout()
    set_frame_size(args.size)
    if not args.frames:
        world = int(os.environ.get('WORLD_SIZE', '1'))
        args.frames = 4096 if FRAME_H * FRAME_W <= 640 * 480 else max(1, 32768 // max(world, args.gpus, 1))
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
