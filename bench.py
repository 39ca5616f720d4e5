#!/usr/bin/env python
"""Throughput of the dial-reading hot path: meter readings/sec on 640x480 frames.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the whole path (locate dials + read 4 dials + value) over one
HBM-resident synthetic batch (4096 frames of 640x480x3 u8 per GPU = 3.77 GB, larger than L2, so no
flush between steps).  Frames are "S-planted" (SURVEY.md 8d): uniform noise canvases with decoded
sample meter crops pasted at the meter_rect with a seeded +-6 px jitter, so every stage runs on every
frame and the expected readings are known.  Params reach the engine the product's own way:
``params.yml`` + template file -> ``_params.load`` -> ``_reading.get_engine``.

Prints ONE JSON line (rank 0).  Besides the headline (BASELINE.json configs[2]) it carries
``config3`` (configs[3]: 1280x720 x 32768 frames in total, strong-scaled over the GPUs), ``set2``
(the second sample geometry), ``micro_bench`` (configs[4], N=1), ``ingest`` (JPEG bytes -> readings,
every N) and ``cpu_baseline`` (the unmodified reference on the host cores + the C port).  Frames shard
by index over the GPUs: no data-path collective; torch.distributed is used for the barrier, the
max-over-ranks time and the host-side gather of records that rank 0 checks for EVERY rank.
"""
import argparse
import json
import math
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))        # fixtures only (golden crops, params as arrays)

METRIC = 'meter readings/sec (640x480 frames)'
UNIT = 'frames/s'
TMPL_W, TMPL_H = 188, 119


def load_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as fp:
            d = json.load(fp)
        return dict(hbm_gbs=d['hbm_gbs'], bf16_tflops=d['bf16_tflops'],
                    bf16_tflops_sustained=d.get('bf16_tflops_sustained', d['bf16_tflops']), source='measured')
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1590.0, source='fallback')


def load_traffic():
    """dram__bytes_read + dram__bytes_write per launch from the committed ncu captures
    (tools/ncu_summary.py traffic ... > profiles/r0N_traffic.json, newest round first)."""
    for name in ('r02_traffic.json', 'r01_traffic.json'):
        path = os.path.join(ROOT, 'profiles', name)
        if os.path.exists(path):
            with open(path) as fp:
                return json.load(fp)
    return {}


class NvmlSampler:
    """SM clock, power and throttle reasons polled in-process through NVML every ~2 ms, so that a
    timed region of a few tens of milliseconds still gets a useful number of samples."""

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
            out.update(sm_mhz=float(np.median(busy)), sm_mhz_min=float(min(busy)), power_w_max=pmax,
                       reasons=sorted(name for (name, bit) in self.REASONS if mask & bit))
        return out


class NullSampler:
    def stop(self):
        return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0, 'sampler': 'unavailable'}


def start_clock_sampler(gpu_index):
    try:
        return NvmlSampler(gpu_index)
    except Exception:
        return NullSampler()


def bind_to_gpu_cpus(gpu_index):
    """Pin this rank to the CPUs NVML lists as local to its GPU before any page-locked buffer is
    allocated (first touch puts the pages on that NUMA node).  Returns what was done."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for (i, w) in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        try:
            numa = int(pynvml.nvmlDeviceGetNumaNodeId(h))
        except Exception:
            numa = None
        return {'cpus': len(cpus), 'cpu_first': min(cpus) if cpus else None,
                'cpu_last': max(cpus) if cpus else None, 'numa_node': numa}
    except Exception as error:                          # noqa: BLE001 - reporting only
        return {'cpus': len(os.sched_getaffinity(0)), 'error': str(error)[:80]}


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


# ---------------------------------------------------------------------------------------------
# workloads: sample params through the product's own setup path, S-planted synthetic frames
# ---------------------------------------------------------------------------------------------
class Workload:
    """Sample params `set_no` (from the committed fixture arrays: the reference tree does not exist
    on the GPU box) written out as params.yml + template file and loaded the way a user's are:
    ``_params.load`` -> ``_reading.get_engine`` (cv2-free dial masks, melf_create)."""

    def __init__(self, set_no, tmpdir, device):
        from helpers import golden_set, write_params_file
        from meterelf_b200 import _params, _reading
        self.gs = golden_set(set_no)
        d = os.path.join(tmpdir, 'set%d' % set_no)
        os.makedirs(d, exist_ok=True)
        self.params_file = write_params_file(self.gs, d)
        self.params = _params.load(self.params_file)
        self.names = list(self.params.dial_centers.keys())
        self.rect = self.params.meter_rect.top_left + self.params.meter_rect.bottom_right
        self.engine = None if device is None else _reading.get_engine(self.params, cropped_input=False, device=device)
        sub = self.gs.subset()
        names = [str(s) for s in sub['names']]
        keep = [i for (i, fn) in enumerate(names) if self.gs.expected[fn]['status'] == 0]
        self.crops = sub['crops'][keep]                    # decoded sample meter_rect crops that read fine

    @property
    def crop_hw(self):
        (x0, y0, x1, y1) = self.rect
        return (y1 - y0, x1 - x0)

    def macs_per_frame(self):
        (ch, cw) = self.crop_hw
        return (cw - TMPL_W + 1) * (ch - TMPL_H + 1) * TMPL_W * TMPL_H


def planted_offsets(n, seed):
    return np.random.default_rng(seed).integers(-6, 7, size=(n, 2))


def make_frames_device(torch, wl, n, fh, fw, seed, device):
    """(n, fh, fw, 3) u8 on the device: noise canvases, sample crops pasted at meter_rect +- 6 px."""
    g = torch.Generator(device=device)
    g.manual_seed(seed + 1)
    frames = torch.empty((n, fh, fw, 3), dtype=torch.uint8, device=device)
    chunk = max(1, (1 << 30) // (fh * fw * 3))
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        frames[a:b] = torch.randint(0, 256, (b - a, fh, fw, 3), dtype=torch.uint8, device=device, generator=g)
    d_crops = torch.from_numpy(np.ascontiguousarray(wl.crops)).to(device)
    (x0, y0, x1, y1) = wl.rect
    off = planted_offsets(n, seed)
    for i in range(n):
        (dx, dy) = (int(off[i][0]), int(off[i][1]))
        frames[i, y0 + dy:y1 + dy, x0 + dx:x1 + dx] = d_crops[i % len(wl.crops)]
    return frames


def make_crops_host(wl, n, seed):
    """What the meter_rect window of make_frames_device-style frames looks like, built on the host
    for the reference arm (noise + the jittered sample crop, cut to meter_rect)."""
    rng = np.random.default_rng(seed + 1)
    (ch, cw) = wl.crop_hw
    pad = 6
    out = np.empty((n, ch, cw, 3), np.uint8)
    off = planted_offsets(n, seed)
    for i in range(n):
        canvas = rng.integers(0, 256, size=(ch + 2 * pad, cw + 2 * pad, 3), dtype=np.uint8)
        (dx, dy) = off[i]
        canvas[pad + dy:pad + dy + ch, pad + dx:pad + dx + cw] = wl.crops[i % len(wl.crops)]
        out[i] = canvas[pad:pad + ch, pad:pad + cw]
    return out


def port_read(wl, crops, n_threads):
    """The C port of the reference (oracle/melf_oracle.c) over meter_rect crops; (records, seconds)."""
    import oracle
    O = wl.gs.oracle_params()
    t0 = time.perf_counter()
    res = oracle.read_batch(O, crops, n_threads=n_threads)
    return (res, time.perf_counter() - t0)


def records_differ(a, b, n_dials):
    return not (a['status'] == b['status'] and a['match_x'] == b['match_x'] and a['match_y'] == b['match_y']
                and a['failed_dial'] == b['failed_dial'] and a['unreadable_mask'] == b['unreadable_mask']
                and np.array_equal(a['dial_pos'][:n_dials], b['dial_pos'][:n_dials], equal_nan=True)
                and (a['value'] == b['value'] or (a['value'] != a['value'] and b['value'] != b['value'])))


class Dist:
    """The little of torch.distributed the bench needs (identity at N=1)."""

    def __init__(self, torch, device):
        import torch.distributed as dist
        self.torch, self.dist, self.device = torch, dist, device
        self.rank = int(os.environ.get('RANK', '0'))
        self.world = int(os.environ.get('WORLD_SIZE', '1'))
        if self.world > 1:
            os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')   # NCCL's banner goes to stdout by default
            dist.init_process_group('nccl', device_id=device)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, value, op='max'):
        t = self.torch.tensor([float(value)], dtype=self.torch.float64, device=self.device)
        if self.world > 1:
            self.dist.all_reduce(t, op={'max': self.dist.ReduceOp.MAX, 'sum': self.dist.ReduceOp.SUM,
                                        'min': self.dist.ReduceOp.MIN}[op])
        return float(t.item())

    def all_values(self, value):
        """One float from every rank, in rank order (on every rank)."""
        t = self.torch.zeros(self.world, dtype=self.torch.float64, device=self.device)
        t[self.rank] = float(value)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [float(v) for v in t.cpu()]

    def gather(self, arr):
        """Rank-ordered concatenation of equal-length host arrays on rank 0 (None elsewhere):
        the product's own host-side gather, ``_sharding.gather_records``."""
        from meterelf_b200 import _sharding
        return _sharding.gather_records(arr, self.world * len(arr), dst=0)

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# one workload on the GPUs: device-resident steps, per-kernel times, end to end, per-rank check
# ---------------------------------------------------------------------------------------------
def measure_workload(torch, D, wl, fh, fw, n, steps, warmup, min_seconds, local, e2e_frames, check_frames,
                     want_kernels=True):
    """n frames of fw x fh per GPU.  Returns (summary dict for rank 0, device frames, last records)."""
    from meterelf_b200 import _engine
    eng = wl.engine
    device = D.device
    d_frames = make_frames_device(torch, wl, n, fh, fw, 1234 + D.rank, device)
    d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device=device)
    stream = torch.cuda.current_stream()
    (row_stride, frame_stride) = (fw * 3, fh * fw * 3)

    def step():
        eng.read_device(d_frames.data_ptr(), n, fh, fw, row_stride, frame_stride, d_out.data_ptr(), stream.cuda_stream)

    def timed(k):
        """k steps between two events, barrier + synchronize on both sides; max over ranks, ms."""
        D.barrier()
        sampler = start_clock_sampler(local) if D.rank == 0 else None
        launches0 = _engine.launch_count()
        (e0, e1) = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        e0.record(stream)
        for _ in range(k):
            step()
        e1.record(stream)
        D.barrier()
        ms = D.reduce(e0.elapsed_time(e1), 'max')
        return (ms, _engine.launch_count() - launches0, sampler.stop() if sampler else None)

    for _ in range(max(warmup, 3)):
        step()
    (burst_ms, burst_launches, burst_clocks) = timed(steps)

    def kernel_times():
        """The same steps with CUDA events around every launch (melf_profile_*): ms per step and kernel kind."""
        ksteps = max(3, min(steps, 20))
        eng.profile_enable(True)
        for _ in range(ksteps):
            step()
        prof = eng.profile_read()
        eng.profile_enable(False)
        return ({k: (v[0] / ksteps) for (k, v) in prof.items() if v[1]},
                {k: v[1] // ksteps for (k, v) in prof.items() if v[1]})
    # once right after the burst (cool part, full clocks: pairs with the burst peak) ...
    burst_kernels = kernel_times()[0] if want_kernels else None
    # the headline is the sustained number: keep stepping until the timed region is min_seconds long
    k2 = max(steps, int(math.ceil(min_seconds * 1e3 / max(burst_ms / steps, 1e-3))))
    k2 = int(D.reduce(k2, 'max'))
    (sus_ms, sus_launches, sus_clocks) = timed(k2)
    res = d_out.cpu().numpy().view(_engine.RESULT_DTYPE).reshape(n).copy()

    out = {
        'frames_per_gpu': n, 'frame_w': fw, 'frame_h': fh, 'steps': k2, 'ms_per_step': sus_ms / k2,
        'value': D.world * n * k2 / (sus_ms / 1e3), 'timed_region_s': sus_ms / 1e3, 'clocks': sus_clocks,
        'gpu_launches': int(sus_launches),
        'burst': {'steps': steps, 'ms_per_step': burst_ms / steps, 'value': D.world * n * steps / (burst_ms / 1e3),
                  'clocks': burst_clocks, 'gpu_launches': int(burst_launches)},
    }

    # ---- ... and once right after the sustained loop (power-capped clocks: pairs with the sustained peak) ----
    if want_kernels:
        out['locate_kernel'] = eng.last_locate()
        (out['kernel_ms_per_step'], out['kernel_launches_per_step']) = kernel_times()
        out['burst']['kernel_ms_per_step'] = burst_kernels

    # ---- end to end through the host API: pinned host frames in, host records out --------------
    if e2e_frames:
        n_e2e = min(n, e2e_frames)
        h_frames = torch.empty((n_e2e, fh, fw, 3), dtype=torch.uint8).pin_memory()
        h_frames.copy_(d_frames[:n_e2e])
        h_np = h_frames.numpy()
        eng.read_frames(h_np[:64])
        esteps = max(2, min(steps, 10))
        D.barrier()
        t0 = time.perf_counter()
        for _ in range(esteps):
            h_res = eng.read_frames(h_np)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        (ch, cw) = wl.crop_hw
        h2d = int(n_e2e * ch * cw * 3)
        rates = D.all_values(h2d * esteps / dt / 1e9)
        out['e2e'] = {'value': D.world * n_e2e * esteps / D.reduce(dt, 'max'), 'unit': UNIT,
                      'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': int(n_e2e * _engine.RESULT_DTYPE.itemsize),
                      'frames_per_step': n_e2e, 'steps': esteps, 'h2d_gbs_per_rank': [round(r, 2) for r in rates],
                      'route': 'melf_read_batch: page-locked host BGR frames, only the meter_rect rows cross PCIe'}
        same = int(np.array_equal(h_res['status'], res['status'][:n_e2e]) and
                   h_res['dial_pos'].tobytes() == res['dial_pos'][:n_e2e].tobytes())
        out['e2e']['records_equal_device_path_all_ranks'] = bool(D.reduce(same, 'min'))
        del h_frames

    # ---- every rank's records against the CPU port on a sample of that rank's own frames -------
    k = min(n, check_frames)
    (x0, y0, x1, y1) = wl.rect
    crops = np.ascontiguousarray(d_frames[:k, y0:y1, x0:x1].cpu().numpy())
    g_crops = D.gather(crops.reshape(k, -1).view(np.dtype((np.void, crops[0].size))).reshape(k))
    g_recs = D.gather(res[:k].copy())
    ok = D.all_values(int(np.sum(res['status'] == 0)))
    if D.rank == 0:
        (ch, cw) = wl.crop_hw
        all_crops = np.frombuffer(g_crops.tobytes(), np.uint8).reshape(D.world * k, ch, cw, 3)
        cores = len(os.sched_getaffinity(0))
        (port, port_dt) = port_read(wl, all_crops, cores)
        mism = [sum(records_differ(g_recs[r * k + i], port[r * k + i], len(wl.names)) for i in range(k))
                for r in range(D.world)]
        out['check'] = {'frames_read_ok_per_rank': [int(v) for v in ok], 'frames_per_rank': n,
                        'checked_per_rank': k, 'mismatches_vs_port_per_rank': mism,
                        'how': 'records and input crops of every rank gathered on rank 0 (_sharding.gather_records)'}
        out['_port'] = {'value': D.world * k / port_dt, 'cores': cores, 'frames': D.world * k}
        out['_crops_all'] = all_crops
        out['_recs_all'] = g_recs
    return (out, d_frames, res)


def ingest_leg(torch, D, wl, n_streams, reps=3):
    """SURVEY 8f-1, at every N: the reference's real input is JPEG files, so this leg times
    melf_read_jpeg_batch -- host JPEG bytes in (page-locked), host records out, decode bit-identical
    to cv2.imread on the GPU -- over the portrait sample files of set 1 repeated to n_streams per GPU,
    and checks every rank's readings against the reference's golden lines."""
    from helpers import GOLDEN, format_result_line
    from test_jpeg_oracle import load_streams
    from meterelf_b200 import _engine
    eng = wl.engine
    full = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    streams = load_streams(full if os.path.exists(full) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
    base = [s for s in streams if s[4] == 1 and s[2] == (640, 480)]
    shift = D.rank * 7                                   # every rank reads the files in another order
    data = [base[(i + shift) % len(base)][1] for i in range(n_streams)]
    (blob, offsets) = _engine.pack_streams(data)
    blob_p = torch.from_numpy(blob.copy()).pin_memory().numpy()
    results = np.zeros((n_streams,), _engine.RESULT_DTYPE)

    def call():
        eng._check(eng.lib.melf_read_jpeg_batch(eng._ctx, blob_p.ctypes.data, offsets.ctypes.data,
                                                n_streams, results.ctypes.data))
    for _ in range(2):
        call()
    launches0 = _engine.launch_count()
    D.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        call()
    dt = (time.perf_counter() - t0) / reps
    launches = (_engine.launch_count() - launches0) // reps
    bad = 0
    for i in range(len(base)):
        exp = wl.gs.expected[base[(i + shift) % len(base)][0]]
        got = format_result_line(results[i], wl.names)
        bad += (got != exp['line']) if exp['status'] == 0 else (
            got.split(' (match val')[0] != exp['line'].split(' (match val')[0])
    rates = D.all_values(float(offsets[-1]) / dt / 1e9)
    bad_all = D.all_values(bad)
    return {'metric': 'meter readings/sec from JPEG bytes (480x640 4:2:0 files, %.1f KB avg)' % (int(offsets[-1]) / n_streams / 1024),
            'value': D.world * n_streams / D.reduce(dt, 'max'), 'unit': UNIT, 'streams_per_gpu': n_streams, 'steps': reps,
            'ms_per_step': D.reduce(dt, 'max') * 1e3, 'h2d_bytes_per_step': int(offsets[-1]),
            'd2h_bytes_per_step': int(results.nbytes), 'gpu_launches_per_step': int(launches),
            'h2d_gbs_per_rank': [round(r, 2) for r in rates],
            'golden_lines_checked_per_rank': len(base), 'golden_line_mismatches_per_rank': [int(b) for b in bad_all],
            'distinct_streams': len(base),
            'workload_note': 'the streams are the %d sample files repeated; the Huffman stage orders a chunk by entropy-coded '
                             'length in 1 KB buckets, so a warp decodes different files, as in real use (an exact sort would '
                             'put byte-identical copies into one warp: profiles/r02_huffman_variants.txt, DESIGN.md section 10)'
                             % len(base),
            'route': 'melf_read_jpeg_batch (the default route of get_meter_values for baseline JPEG files): header walk '
                     'on the host, Huffman / IDCT / upsampling / colour on the GPU (bit-identical to cv2.imread), '
                     'then the same locate and dial kernels; GPU-bound (Huffman stage), DESIGN.md section 8'}


def api_leg(wl, n_files):
    """The drop-in call itself: get_meter_values(params_file, filenames) over JPEG FILES (the sample files
    of set 1 written to a RAM disk, repeated), file reads, header probes, GPU decode + read and the Python
    objects of every result included (reference: _api.py:16-33, one cv2.imread + get_meter_value per file)."""
    import shutil
    from helpers import GOLDEN
    from test_jpeg_oracle import load_streams
    import meterelf_b200
    full = os.path.join(GOLDEN, '_full', 'jpeg_all.npz')
    streams = load_streams(full if os.path.exists(full) else os.path.join(GOLDEN, 'jpeg_subset.npz'))
    base = [s for s in streams if s[4] == 1 and s[2] == (640, 480)]
    root = '/dev/shm' if os.path.isdir('/dev/shm') and os.access('/dev/shm', os.W_OK) else None
    d = tempfile.mkdtemp(prefix='melf_api_', dir=root)
    try:
        paths = []
        for (k, (name, data, _shape, _sha, _set)) in enumerate(base):
            paths.append(os.path.join(d, name))
            with open(paths[-1], 'wb') as fp:
                fp.write(data)
        names = [paths[i % len(paths)] for i in range(n_files)]
        list(meterelf_b200.get_meter_values(wl.params_file, names))          # warm-up (page-locked blob pool at its final size)
        runs = []
        for _ in range(3):          # host-bound and noisy from box to box and run to run: the median of three
            t0 = time.perf_counter()
            got = list(meterelf_b200.get_meter_values(wl.params_file, names))
            runs.append(time.perf_counter() - t0)
        dt = sorted(runs)[1]
        bad = 0
        for (i, item) in enumerate(got[:len(base)]):
            exp = wl.gs.expected[base[i][0]]
            line = '{:07.3f}'.format(item.value) if item.value else ('UNKNOWN ' + item.error.get_message() if item.error else '')
            bad += (line != exp['line']) if exp['status'] == 0 else (
                line.split(' (match val')[0] != exp['line'].split(' (match val')[0])
    finally:
        shutil.rmtree(d, ignore_errors=True)
    return {'metric': 'get_meter_values(params_file, filenames) over JPEG files on a RAM disk', 'value': n_files / dt,
            'unit': 'files/s', 'files': n_files, 'seconds': dt, 'files_per_s_all_runs': [round(n_files / t) for t in runs],
            'golden_line_mismatches': int(bad),
            'golden_lines_checked': len(base),
            'note': 'one Python process: file reads and header probes batched on host threads behind the C ABI into '
                    'pooled page-locked buffers, two batches in flight (files of batch k + 1 read under the GPU work of '
                    'batch k), results converted once per batch; bound by the host (file system calls, Python objects), '
                    'not the GPU'}


def convert_stage(torch, wl, d_frames, fh, fw, steps, local, peaks, traffic):
    """Stage 1 as a standalone fused pass over full frames (north_star's >= 50 % of HBM target)."""
    from meterelf_b200 import _engine
    nconv = min(len(d_frames), 1024)
    device = d_frames.device
    stream = torch.cuda.current_stream()
    d_l = torch.empty((nconv, fh, fw), dtype=torch.uint8, device=device)
    d_m = torch.empty((nconv, fh, fw), dtype=torch.uint8, device=device)

    def convert_only():
        _engine.convert_hls_device(local, d_frames.data_ptr(), nconv, fh, fw, fw * 3, fh * fw * 3,
                                   wl.gs.hue_shift, True, 0, d_l.data_ptr(),
                                   (116, 35, 95), (134, 125, 165), d_m.data_ptr(), stream.cuda_stream)
    for _ in range(3):
        convert_only()
    torch.cuda.synchronize()
    csteps = max(5, min(steps, 50))
    (c0, c1) = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
    c0.record(stream)
    for _ in range(csteps):
        convert_only()
    c1.record(stream)
    torch.cuda.synchronize()
    conv_ms = c0.elapsed_time(c1) / csteps
    gbs = 5.0 * fh * fw * nconv / (conv_ms / 1e3) / 1e9
    return {'kernel': 'melf_convert_tab_kernel (fused BGR->HLS + hue shift + L plane + mask, full frames)',
            'bound': 'hbm', 'achieved': gbs, 'peak': peaks['hbm_gbs'], 'unit': 'GB/s', 'frac': gbs / peaks['hbm_gbs'],
            'traffic': traffic.get('melf_convert_tab_kernel'), 'ms_per_launch': conv_ms,
            'note': 'algorithmic 5 B/px x %d px; peak %s' % (fh * fw * nconv, peaks['source'])}


def micro_bench(torch, wl, local, int8_peak):
    """BASELINE.json configs[4]: the 188x119 dial template against single-channel search windows of
    growing size through melf_match_template (cv2.matchTemplate TM_CCOEFF + minMaxLoc semantics);
    the small windows are checked against exact numpy int64 arithmetic."""
    import oracle
    eng = wl.engine
    tmpl = np.ascontiguousarray(wl.gs.template)
    (th, tw) = tmpl.shape
    dev = torch.device('cuda', local)
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    rows = []
    st = torch.cuda.current_stream()
    for (w, h) in [(220, 135), (250, 250), (320, 240), (640, 480), (1280, 720)]:
        n = max(8, min(4096, (2048 << 20) // (w * h)))
        frames = torch.randint(0, 256, (n, h, w), dtype=torch.uint8, device=dev, generator=g)
        frames[0, 3:3 + th, 5:5 + tw] = torch.from_numpy(tmpl).to(dev)
        d_xy = torch.zeros((n, 2), dtype=torch.int32, device=dev)
        d_val = torch.zeros((n,), dtype=torch.float32, device=dev)

        def run():
            eng.match_template_device(frames.data_ptr(), 1, n, h, w, w, h * w, False,
                                      d_xy.data_ptr(), d_val.data_ptr(), 0, st.cuda_stream)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        ok = (int(d_xy[0, 0]), int(d_xy[0, 1])) == (5, 3)
        if w * h <= 320 * 240:                          # the CPU port's exact integer arg-max
            (mx, my, mv, _map) = oracle.match_template(frames[1].cpu().numpy(), tmpl, want_map=False)
            ok = ok and (int(d_xy[1, 0]), int(d_xy[1, 1]), float(d_val[1])) == (mx, my, float(mv))
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
        rows.append({'window': '%dx%d' % (w, h), 'result_map': '%dx%d' % (w - tw + 1, h - th + 1), 'frames': n,
                     'ms_per_batch': ms, 'frames_per_s': n / (ms / 1e3), 'algorithmic_tops': tops,
                     'frac_of_int8_peak': tops / int8_peak, 'argmax_check': bool(ok)})
        del frames
    return {'what': 'template-correlation micro-bench: 188x119 template, window-size sweep (configs[4])',
            'int8_peak_tops': int8_peak, 'rows': rows}


def sharded_check(D, wl):
    """The product's multi-GPU path with the real Engine on every GPU: the SAME global batch on every
    rank, _sharding.read_frames_sharded reads the rank's slice and gathers; rank 0 compares all of it
    with the CPU port."""
    from meterelf_b200 import _reading, _sharding
    eng = _reading.get_engine(wl.params, cropped_input=True, device=D.device.index)
    crops = np.ascontiguousarray(np.concatenate([wl.crops] * 4)[:16 * D.world + 3])
    got = _sharding.read_frames_sharded(eng.read_frames, crops, dst=0)
    if D.rank != 0:
        return None
    (want, _dt) = port_read(wl, crops, 4)
    bounds = _sharding.shard_bounds(len(crops), D.world)
    return {'frames': len(crops), 'how': '_sharding.read_frames_sharded(Engine.read_frames, frames), gather on rank 0',
            'mismatches_vs_port_per_rank': [
                sum(records_differ(got[i], want[i], len(wl.names)) for i in range(a, b)) for (a, b) in bounds]}


def reference_pool_run(wl, crops, cores, runs=1):
    """The UNMODIFIED reference (oracle/_ref, one process per core) over meter_rect crops:
    (frames/s, records) or (None, None) when oracle/_ref is absent."""
    import oracle
    if not oracle.ref_available():
        return (None, None)
    with oracle.RefPool(wl.params_file, crops, cores) as pool:
        walls = []
        for _ in range(runs):
            (wall, recs) = pool.run()
            walls.append(wall)
    return (len(crops) / (sum(walls) / len(walls)), recs)


def workload_config(wl, fw, fh, n, world, total=None):
    (ch, cw) = wl.crop_hw
    return {
        'workload': 'synthetic %dx%d (w x h) BGR batch of %d frames per GPU%s, S-planted sample meter crops '
                    '(sample-images%d params, meter_rect %dx%d, template %dx%d)'
                    % (fw, fh, n, (' (%d in total, strong-scaled)' % total) if total else '', wl.gs.set_no, cw, ch,
                       TMPL_W, TMPL_H),
        'frames_per_gpu': n, 'frame_h': fh, 'frame_w': fw, 'bytes_per_frame': fh * fw * 3,
        'l2': 'batch (%.2f GB per GPU) is larger than L2; no flush between steps' % (n * fh * fw * 3 / 1e9),
        'parallelism': 'frames sharded by index, one process per GPU (%d), host-side gather of records' % world,
        'params': 'params.yml + template file -> _params.load -> _reading.get_engine',
    }


def locate_roofline(wl, km, n, step_ms, peaks, traffic, km_burst=None, name='melf_match_fused_ov_kernel'):
    """Roofline object of the dominant kernel (tensor-bound).  `achieved` / `frac`: the kernel
    timed right after the burst against the burst peak; `sustained`: timed right after the seconds-long loop
    (power-capped clocks) against the sustained peak."""
    tc_sus = km.get('match_tc')
    tc_ms = (km_burst or km).get('match_tc')
    if not tc_ms:
        return None
    macs = wl.macs_per_frame()
    tops = 2.0 * macs * n / (tc_ms / 1e3) / 1e12
    tops_sus = 2.0 * macs * n / (tc_sus / 1e3) / 1e12 if tc_sus else None
    peak_burst = 2.0 * peaks['bf16_tflops']
    peak_sus = 2.0 * peaks['bf16_tflops_sustained']
    issue = 8192 * 2 * 148 * 1.965e9 / 1e12
    (ch, cw) = wl.crop_hw
    crop_bytes = ch * cw * 3 * n
    return {
        'kernel': name + (' (+ melf_small_edge_kernel; positions as M, 32 frames as N, BGR crop -> lightness in the shadow of '
                          'the MMAs; TM_CCOEFF on tcgen05 kind::i8, exact integer correlation + arg-max)'
                          if name == 'melf_small_fused_kernel' else
                          ' (BGR crop -> lightness + window sums in the shadow of the MMAs, drain under the next '
                          'pass; TM_CCOEFF on tcgen05 kind::i8, exact integer correlation + arg-max)'),
        'bound': 'tensor', 'achieved': tops, 'peak': peak_burst, 'unit': 'TOP/s', 'frac': tops / peak_burst,
        'traffic': traffic.get(name), 'ms_per_launch': tc_ms,
        'sustained': {'achieved': tops_sus, 'peak': peak_sus, 'frac': tops_sus / peak_sus if tops_sus else None,
                      'ms_per_launch': tc_sus},
        'peak_int8_issue_rate': issue, 'frac_of_int8_issue_rate': tops / issue,
        'hbm_view': {'algorithmic_bytes': crop_bytes, 'achieved_gbs': crop_bytes / (tc_ms / 1e3) / 1e9,
                     'frac_of_hbm_peak': crop_bytes / (tc_ms / 1e3) / 1e9 / peaks['hbm_gbs']},
        'note': 'algorithmic 2*%d op/frame x %d frames per launch (SURVEY 8d: result positions x 22372 MAC); '
                'ms_per_launch = CUDA events around the launch on its own stream: achieved / frac = timed right after '
                'the 20-step burst against peak = 2 x %s cuBLAS bf16 dense burst (int8:bf16 tensor ratio), sustained = '
                'timed right after the seconds-long loop (power cap) against 2 x the sustained cuBLAS figure; '
                'peak_int8_issue_rate = 8192 MAC/clk/SM x 148 SMs x 1965 MHz; kernel share of the sustained step %.2f'
                % (macs, n, peaks['source'], (tc_sus or tc_ms) / step_ms),
    }

def run_ours(args):
    import torch
    from meterelf_b200 import _engine

    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise _engine.EngineUnavailableError('bench needs a CUDA device; there is no CPU fallback')
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    affinity = bind_to_gpu_cpus(local)
    D = Dist(torch, device)
    tmp = tempfile.TemporaryDirectory()
    peaks = load_peaks()
    traffic = load_traffic()
    wl = Workload(1, tmp.name, local)
    (fh, fw, n) = (640, 480, args.frames)      # the sample frames are portrait: 480 wide x 640 tall

    (head, d_frames, res) = measure_workload(torch, D, wl, fh, fw, n, args.steps, args.warmup, args.min_seconds,
                                             local, args.e2e_frames, args.check_frames)
    stage = [convert_stage(torch, wl, d_frames, fh, fw, args.steps, local, peaks, traffic)]

    # ---- CPU baseline on rank 0: the unmodified reference, one process per core, + the C port ------
    cpu = None
    if D.rank == 0:
        import oracle
        cores = len(os.sched_getaffinity(0))
        k = head['check']['checked_per_rank']
        sample = min(n, max(cores * args.cpu_frames_per_core, k))
        (x0, y0, x1, y1) = wl.rect
        crops0 = np.ascontiguousarray(d_frames[:sample, y0:y1, x0:x1].cpu().numpy())
        crops_all = np.concatenate([crops0, head['_crops_all'][k:]])         # + the other ranks' check frames
        recs_all = np.concatenate([res[:sample], head['_recs_all'][k:]])
        (ref_fps, ref_recs) = reference_pool_run(wl, crops_all, cores)
        (port, port_dt) = port_read(wl, crops_all, cores)
        port_fps = len(crops_all) / port_dt
        cpu = {'value': ref_fps if ref_fps else port_fps, 'unit': UNIT, 'cores': cores,
               'kind': 'reference' if ref_fps else 'port',
               'sample': '%d frames: the first %d of rank 0\'s batch + %d check frames of every other rank'
                         % (len(crops_all), sample, k),
               'what': 'unmodified reference package (oracle/_ref), ImageFile(name, params, bgr_image=crop) -> '
                       'get_meter_value, one process per core, cv2 single-threaded' if ref_fps else
                       'oracle/_ref absent: C port of the reference (oracle/melf_oracle.c), one thread per core',
               'port': {'value': port_fps, 'unit': UNIT, 'cores': cores, 'kind': 'port'},
               'mismatches_vs_gpu': (oracle.ref_mismatches(ref_recs, recs_all, len(wl.names)) if ref_recs else
                                     sum(records_differ(a, b, len(wl.names)) for (a, b) in zip(recs_all, port))),
               'port_mismatches_vs_gpu': sum(records_differ(a, b, len(wl.names)) for (a, b) in zip(recs_all, port))}
    del d_frames
    torch.cuda.empty_cache()

    ingest = ingest_leg(torch, D, wl, args.ingest_streams) if args.ingest_streams > 0 else None
    shard = sharded_check(D, wl)

    # ---- BASELINE.json configs[3]: 1280x720, 32768 frames in total over the GPUs (strong scaling) ----
    c3 = None
    if args.config3_frames > 0:
        n3 = max(1, args.config3_frames // D.world)
        (c3, d3, _r3) = measure_workload(torch, D, wl, 720, 1280, n3, max(3, min(args.steps, 10)), 3,
                                         min(args.min_seconds, 1.0), local, min(args.e2e_frames, 2048), 32)
        del d3
        torch.cuda.empty_cache()

    # ---- the second sample geometry (220x135 meter_rect, 33x17 result map) ---------------------------
    wl2 = Workload(2, tmp.name, local)
    (s2, d2, _r2) = measure_workload(torch, D, wl2, fh, fw, n, args.steps, 3, min(args.min_seconds, 0.5), local, 0, 64)
    del d2
    torch.cuda.empty_cache()

    if D.rank != 0:
        D.close()
        return

    def public(d):
        return {k: v for (k, v) in d.items() if not k.startswith('_')}

    km = head['kernel_ms_per_step']
    step_ms = head['ms_per_step']
    roofline = locate_roofline(wl, km, n, step_ms, peaks, traffic, head['burst'].get('kernel_ms_per_step'),
                               head.get('locate_kernel') or 'melf_match_fused_ov_kernel')
    if km.get('dials'):
        stage.append({'kernel': 'melf_dials_kernel (per-dial masks, contour selection, angles, value)',
                      'bound': 'issue/latency (SURVEY 8d stage 3)', 'ms_per_launch': km['dials'],
                      'frames_per_s': n / (km['dials'] / 1e3), 'traffic': traffic.get('melf_dials_kernel'),
                      'note': 'timed alone; share of the sum of kernel times %.2f' % (km['dials'] / sum(km.values()))})
    int8_peak = 2.0 * peaks['bf16_tflops']
    line = {
        'metric': METRIC, 'value': head['value'], 'unit': UNIT, 'n_gpus': D.world, 'steps': head['steps'],
        'steps_requested': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': step_ms,
        'timed_region_s': head['timed_region_s'], 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'u8/int32 (correlation), f32 (HLS), f64 (angles)', 'data': 'synthetic',
        'config': workload_config(wl, fw, fh, n, D.world),
        'e2e': head.get('e2e'), 'gpu_launches': head['gpu_launches'], 'clocks': head['clocks'],
        'burst': head['burst'], 'roofline': roofline, 'stage_rooflines': stage,
        'kernel_ms_per_step': km, 'kernel_launches_per_step': head['kernel_launches_per_step'],
        'cpu_baseline': cpu, 'check': head['check'], 'sharded_check': shard, 'ingest': ingest,
        'affinity': affinity,
        'note': 'value = sustained: the timed region is max(--steps, --min-seconds worth of) steps; burst = exactly '
                '--steps steps, timed first',
    }
    if c3:
        c3p = public(c3)
        c3p['config'] = workload_config(wl, 1280, 720, c3['frames_per_gpu'], D.world, args.config3_frames)
        c3p.update(metric='meter readings/sec (1280x720 frames)', unit=UNIT, scaling='strong')
        c3p['roofline'] = locate_roofline(wl, c3['kernel_ms_per_step'], c3['frames_per_gpu'], c3['ms_per_step'],
                                          peaks, traffic, c3['burst'].get('kernel_ms_per_step'),
                                          c3.get('locate_kernel') or 'melf_match_fused_ov_kernel')
        line['config3'] = c3p
    s2p = public(s2)
    s2p['config'] = workload_config(wl2, fw, fh, n, D.world)
    s2p['roofline'] = locate_roofline(wl2, s2['kernel_ms_per_step'], n, s2['ms_per_step'], peaks, traffic,
                                      s2['burst'].get('kernel_ms_per_step'), s2.get('locate_kernel') or 'melf_match_fused_ov_kernel')
    line['set2'] = s2p
    if D.world == 1 and args.micro_bench:
        line['micro_bench'] = micro_bench(torch, wl, local, int8_peak)
    if D.world == 1 and args.api_files > 0:
        line['api'] = api_leg(wl, args.api_files)
    emit(line)
    D.close()
    tmp.cleanup()


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on all host cores -- the
    unmodified reference package from oracle/_ref (vendored by __graft_entry__.build()), one process
    per core, each step a bounded sample of the workload (cores x 64 meter_rect crops of S-planted
    frames).  Falls back to the C port only when oracle/_ref is absent."""
    if int(os.environ.get('RANK', '0')) != 0:
        return
    import oracle
    tmp = tempfile.TemporaryDirectory()
    wl = Workload(1, tmp.name, None)
    cores = len(os.sched_getaffinity(0))
    sample = max(cores * args.cpu_frames_per_core, 64)
    crops = make_crops_host(wl, sample, seed=1234)
    times = []
    kind = 'reference' if oracle.ref_available() else 'port'
    if kind == 'reference':
        with oracle.RefPool(wl.params_file, crops, cores) as pool:
            for i in range(args.warmup + args.steps):
                (wall, recs) = pool.run(want_results=(i == 0))
                if i == 0:
                    (port, _dt) = port_read(wl, crops, cores)
                    bad = oracle.ref_mismatches(recs, port, len(wl.names))
                if i >= args.warmup:
                    times.append(wall)
    else:
        bad = None
        for i in range(args.warmup + args.steps):
            (_res, dt) = port_read(wl, crops, cores)
            if i >= args.warmup:
                times.append(dt)
    total = sum(times)
    value = sample * len(times) / total
    cfg = workload_config(wl, 480, 640, args.frames, args.gpus)
    cfg['params'] = 'params.yml + template file -> the reference\'s own _params.load'
    cfg['timed_per_step'] = '%d frames (cores x %d), the meter_rect crops of such frames' % (sample, args.cpu_frames_per_core)
    emit({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * total / len(times),
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8/f32/f64 (OpenCV)',
        'data': 'synthetic', 'config': cfg,
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': kind,
                         'sample': '%d frames per step' % sample, 'port_mismatches_vs_reference': bad},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    })
    tmp.cleanup()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--min-seconds', type=float, default=2.0,
                    help='the headline loops until its timed region is this long (sustained clocks)')
    ap.add_argument('--frames', type=int, default=4096, help='640x480 frames per GPU')
    ap.add_argument('--config3-frames', type=int, default=32768, help='1280x720 frames in total (0 = skip)')
    ap.add_argument('--e2e-frames', type=int, default=4096)
    ap.add_argument('--check-frames', type=int, default=128, help='frames of every rank checked on rank 0')
    ap.add_argument('--cpu-frames-per-core', type=int, default=64, help='CPU baseline sample per host core')
    ap.add_argument('--ingest-streams', type=int, default=32768, help='JPEG streams per GPU and step (0 = skip)')
    ap.add_argument('--micro-bench', type=int, default=1)
    ap.add_argument('--api-files', type=int, default=32768, help='files of the get_meter_values leg (N=1; 0 = skip)')
    args = ap.parse_args()
    protect_stdout()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
