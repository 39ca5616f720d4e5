"""ctypes front-end of the CPU oracle (``oracle/melf_oracle.c``).

TEST INFRASTRUCTURE ONLY: imported by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``; never by the product
package ``meterelf_b200``.
"""
import ctypes as C
import os
import subprocess
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, '_build', 'libmelf_oracle.so')
_JPEG_LIB_PATH = os.path.join(_HERE, '_build', 'libjpeg_oracle.so')
MAX_DIALS = 8


class _Dial(C.Structure):
    _fields_ = [('cx', C.c_double), ('cy', C.c_double), ('angle_of_zero', C.c_double),
                ('range_h', C.c_int32), ('range_l', C.c_int32), ('range_s', C.c_int32),
                ('negative_momentum', C.c_int32),
                ('mask', C.c_void_p), ('circle_mask', C.c_void_p)]


class _Params(C.Structure):
    _fields_ = [('n_dials', C.c_int32), ('tmpl_w', C.c_int32), ('tmpl_h', C.c_int32),
                ('tmpl', C.c_void_p), ('hue_shift', C.c_int32), ('s_tail', C.c_int32),
                ('match_threshold', C.c_double), ('value_order', C.c_int32 * 4),
                ('dials', _Dial * MAX_DIALS)]


class _Result(C.Structure):
    _fields_ = [('status', C.c_int32), ('match_x', C.c_int32), ('match_y', C.c_int32),
                ('max_val', C.c_float), ('failed_dial', C.c_int32),
                ('unreadable_mask', C.c_uint32),
                ('dial_pos', C.c_double * MAX_DIALS), ('value', C.c_double)]


class _Taps(C.Structure):
    _fields_ = [('mask_orig', C.c_void_p), ('mask_de', C.c_void_p),
                ('needle_mask', C.c_void_p), ('dial_color', C.c_void_p),
                ('momentum', C.c_void_p), ('contour_area', C.c_void_p),
                ('n_needle_points', C.c_void_p), ('n_outer_kept', C.c_void_p)]


RESULT_DTYPE = np.dtype([
    ('status', np.int32), ('match_x', np.int32), ('match_y', np.int32),
    ('max_val', np.float32), ('failed_dial', np.int32), ('unreadable_mask', np.uint32),
    ('dial_pos', np.float64, (MAX_DIALS,)), ('value', np.float64)], align=True)

_lib = None


def _stale(lib_path: str, src: str) -> bool:
    return not os.path.exists(lib_path) or (
        os.path.getmtime(lib_path) < os.path.getmtime(os.path.join(_HERE, src)))


def build(force: bool = False) -> str:
    if force or _stale(_LIB_PATH, 'melf_oracle.c') or _stale(_JPEG_LIB_PATH, 'jpeg_oracle.c'):
        subprocess.check_call(['make', '-C', _HERE, '-s'] + (['-B'] if force else []))
    return _LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        assert _lib.melf_o_sizeof_params() == C.sizeof(_Params)
        assert _lib.melf_o_sizeof_result() == C.sizeof(_Result) == RESULT_DTYPE.itemsize
        _lib.melf_o_angle_by_vector.restype = C.c_double
        _lib.melf_o_angle_by_vector.argtypes = [C.c_double, C.c_double]
        _lib.melf_o_value_by_positions.restype = C.c_double
        _lib.melf_o_value_by_positions.argtypes = [C.c_double] * 4
    return _lib


def _ptr(a: Optional[np.ndarray]) -> Optional[int]:
    return None if a is None else a.ctypes.data


class OracleParams:
    """Plain-data view of what the path needs from params.yml + dial masks.

    ``dials``: list (YAML order) of dicts with keys name, center (x, y), color_range
    (h, l, s), negative_momentum, angle_of_zero, mask, circle_mask.
    """

    def __init__(self, template: np.ndarray, hue_shift: int, match_threshold: float,
                 dials: Sequence[Dict[str, Any]], s_tail: bool = True) -> None:
        assert template.dtype == np.uint8 and template.ndim == 2
        assert 1 <= len(dials) <= MAX_DIALS
        self.template = np.ascontiguousarray(template)
        self.names: List[str] = [d['name'] for d in dials]
        self._keep = []
        p = _Params()
        p.n_dials = len(dials)
        (p.tmpl_h, p.tmpl_w) = self.template.shape
        p.tmpl = self.template.ctypes.data
        p.hue_shift = int(hue_shift)
        p.s_tail = 1 if s_tail else 0
        p.match_threshold = float(match_threshold)
        order = sorted(range(len(dials)), key=lambda i: self.names[i])
        for (k, i) in enumerate(order[:4]):
            p.value_order[k] = i
        for (i, d) in enumerate(dials):
            mask = np.ascontiguousarray(d['mask'], dtype=np.uint8)
            ring = np.ascontiguousarray(d['circle_mask'], dtype=np.uint8)
            assert mask.shape == self.template.shape == ring.shape
            self._keep += [mask, ring]
            q = p.dials[i]
            (q.cx, q.cy) = (float(d['center'][0]), float(d['center'][1]))
            q.angle_of_zero = float(d['angle_of_zero'])
            (q.range_h, q.range_l, q.range_s) = (int(v) for v in d['color_range'])
            q.negative_momentum = 1 if d['negative_momentum'] else 0
            q.mask = mask.ctypes.data
            q.circle_mask = ring.ctypes.data
        self.c = p
        self.n_dials = len(dials)
        self.tmpl_h, self.tmpl_w = self.template.shape


def convert_to_hls(bgr: np.ndarray, hue_shift: int = 0, s_tail: bool = True) -> np.ndarray:
    """_utils.py:100-102 on an (h, w, 3) u8 array (any row stride)."""
    assert bgr.dtype == np.uint8 and bgr.ndim == 3 and bgr.shape[2] == 3
    assert bgr.strides[2] == 1 and bgr.strides[1] == 3
    (h, w) = bgr.shape[:2]
    out = np.empty((h, w, 3), np.uint8)
    lib().melf_o_convert_to_hls(C.c_void_p(bgr.ctypes.data), h, w, C.c_ssize_t(bgr.strides[0]),
                                int(hue_shift), 1 if s_tail else 0, C.c_void_p(out.ctypes.data))
    return out


def match_template(img: np.ndarray, tmpl: np.ndarray, want_map: bool = True):
    """_utils.py:91-97; returns (x, y, max_val float32, result map or None)."""
    assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
    tmpl = np.ascontiguousarray(tmpl)
    (h, w) = img.shape
    (th, tw) = tmpl.shape
    res = np.empty((h - th + 1, w - tw + 1), np.float32) if want_map else None
    mx = C.c_int32()
    my = C.c_int32()
    mv = C.c_float()
    lib().melf_o_match_template(
        C.c_void_p(img.ctypes.data), h, w, C.c_ssize_t(img.strides[0]),
        C.c_void_p(tmpl.ctypes.data), th, tw, C.c_void_p(_ptr(res)),
        C.byref(mx), C.byref(my), C.byref(mv))
    return (mx.value, my.value, np.float32(mv.value), res)


def morph3(mask: np.ndarray, erode: bool) -> np.ndarray:
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    out = np.empty_like(mask)
    lib().melf_o_morph3(C.c_void_p(mask.ctypes.data), mask.shape[0], mask.shape[1],
                        1 if erode else 0, C.c_void_p(out.ctypes.data))
    return out


def select_needle_mask(mask_de: np.ndarray, dial_mask: np.ndarray):
    """_reading.py:132-148; returns (have_contours, needle_mask, selected_area)."""
    mask_de = np.ascontiguousarray(mask_de, dtype=np.uint8)
    dial_mask = np.ascontiguousarray(dial_mask, dtype=np.uint8)
    out = np.zeros_like(mask_de)
    area = C.c_double(-1.0)
    have = lib().melf_o_select_needle_mask(
        C.c_void_p(mask_de.ctypes.data), C.c_void_p(dial_mask.ctypes.data),
        mask_de.shape[0], mask_de.shape[1], C.c_void_p(out.ctypes.data), C.byref(area))
    return (bool(have), out, area.value)


def angle_by_vector(x: float, y: float) -> Optional[float]:
    v = lib().melf_o_angle_by_vector(float(x), float(y))
    return None if v != v else v


def value_by_positions(r4: float, r3: float, r2: float, r1: float) -> float:
    return lib().melf_o_value_by_positions(r4, r3, r2, r1)


class Taps:
    def __init__(self, params: OracleParams) -> None:
        (n, h, w) = (params.n_dials, params.tmpl_h, params.tmpl_w)
        self.mask_orig = np.zeros((n, h, w), np.uint8)
        self.mask_de = np.zeros((n, h, w), np.uint8)
        self.needle_mask = np.zeros((n, h, w), np.uint8)
        self.dial_color = np.zeros((n, 3), np.int32)
        self.momentum = np.full((n, 3), np.nan)
        self.contour_area = np.full((n,), -1.0)
        self.n_needle_points = np.zeros((n,), np.int32)
        self.n_outer_kept = np.zeros((n,), np.int32)
        self.hls = None
        self.c = _Taps(*(a.ctypes.data for a in (
            self.mask_orig, self.mask_de, self.needle_mask, self.dial_color, self.momentum,
            self.contour_area, self.n_needle_points, self.n_outer_kept)))


def read_frame(params: OracleParams, bgr: np.ndarray, want_taps: bool = False):
    """Whole path for one meter_rect-cropped BGR frame; returns (result record, taps)."""
    assert bgr.dtype == np.uint8 and bgr.ndim == 3 and bgr.shape[2] == 3
    assert bgr.strides[2] == 1 and bgr.strides[1] == 3
    (h, w) = bgr.shape[:2]
    res = np.zeros((), RESULT_DTYPE)
    taps = Taps(params) if want_taps else None
    hls = np.empty((h, w, 3), np.uint8) if want_taps else None
    rc = lib().melf_o_read_frame(
        C.byref(params.c), C.c_void_p(bgr.ctypes.data), h, w, C.c_ssize_t(bgr.strides[0]),
        C.c_void_p(res.ctypes.data), C.byref(taps.c) if taps else None, C.c_void_p(_ptr(hls)))
    if rc:
        raise ValueError('frame smaller than the template')
    if taps:
        taps.hls = hls
    return (res, taps)


def read_dials_hls(params: OracleParams, dials_hls: np.ndarray, want_taps: bool = False):
    """get_meter_value on a ready dials_hls rect (tmpl_h, tmpl_w, 3)."""
    assert dials_hls.shape == (params.tmpl_h, params.tmpl_w, 3)
    assert dials_hls.strides[2] == 1 and dials_hls.strides[1] == 3
    res = np.zeros((), RESULT_DTYPE)
    taps = Taps(params) if want_taps else None
    lib().melf_o_read_dials_hls(
        C.byref(params.c), C.c_void_p(dials_hls.ctypes.data), C.c_ssize_t(dials_hls.strides[0]),
        C.c_void_p(res.ctypes.data), C.byref(taps.c) if taps else None)
    return (res, taps)


def read_batch(params: OracleParams, frames: np.ndarray, n_threads: int = 1) -> np.ndarray:
    """frames: (n, h, w, 3) u8, C-contiguous in the last two axes."""
    assert frames.dtype == np.uint8 and frames.ndim == 4 and frames.shape[3] == 3
    assert frames.strides[3] == 1 and frames.strides[2] == 3
    (n, h, w) = frames.shape[:3]
    res = np.zeros((n,), RESULT_DTYPE)
    rc = lib().melf_o_read_batch(
        C.byref(params.c), C.c_void_p(frames.ctypes.data), n, h, w,
        C.c_ssize_t(frames.strides[1]), C.c_ssize_t(frames.strides[0]),
        C.c_void_p(res.ctypes.data), int(n_threads))
    if rc:
        raise ValueError('frame smaller than the template')
    return res


def average_translated(frames: np.ndarray, match_xy: np.ndarray, anchor=(30, 116)) -> np.ndarray:
    """CPU restatement (numpy, test infrastructure) of ``get_average_meter_image``
    (``_calibration.py:63-66``): ``get_bgr_image_t``'s integer translation with zero border
    (``_image.py:34-44``; cv2.warpAffine of an integer shift copies pixels exactly),
    ``normalize_image`` (``_utils.py:64-65``), the running mean of ``_image_avg_reducer``
    (``_utils.py:83-88``: ``prev * ((n - 1) / n) + img / n``) and ``denormalize_image``
    (``_utils.py:68-69``)."""
    frames = np.asarray(frames)
    (n, h, w) = frames.shape[:3]
    avg = None
    for f in range(n):
        (tx, ty) = (int(anchor[0]) - int(match_xy[f][0]), int(anchor[1]) - int(match_xy[f][1]))
        shifted = np.zeros_like(frames[f])
        (dx0, dx1) = (max(0, tx), min(w, w + tx))
        (dy0, dy1) = (max(0, ty), min(h, h + ty))
        if dx0 < dx1 and dy0 < dy1:
            shifted[dy0:dy1, dx0:dx1] = frames[f][dy0 - ty:dy1 - ty, dx0 - tx:dx1 - tx]
        norm = shifted.astype(np.float64) / 255.0
        if avg is None:
            avg = norm
        else:
            cnt = f + 1
            avg = avg * ((cnt - 1) / cnt) + (norm / cnt)
    return ((avg * 255.0) + 0.5).astype(np.uint8)


# ---- ingest row: JPEG -> BGR as cv2.imread returns it (oracle/jpeg_oracle.c) ----
_jpeg_lib = None


def _jpeg() -> C.CDLL:
    global _jpeg_lib
    if _jpeg_lib is None:
        build()
        _jpeg_lib = C.CDLL(_JPEG_LIB_PATH)
        _jpeg_lib.jo_decode_bgr.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                            C.POINTER(C.c_int), C.POINTER(C.c_int)]
        _jpeg_lib.jo_decode_coefficients.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_void_p,
                                                     C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    return _jpeg_lib


class JpegError(ValueError):
    pass


def decode_jpeg(data: bytes) -> np.ndarray:
    """``cv2.imread`` of a baseline JPEG held in memory: (h, w, 3) uint8 BGR (_image.py:49)."""
    data = bytes(data)
    (w, h) = (C.c_int(), C.c_int())
    rc = _jpeg().jo_decode_bgr(data, len(data), None, 0, w, h)
    if rc:
        raise JpegError('jpeg oracle: code %d' % rc)
    out = np.empty((h.value, w.value, 3), np.uint8)
    rc = _jpeg().jo_decode_bgr(data, len(data), out.ctypes.data, out.size, w, h)
    if rc:
        raise JpegError('jpeg oracle: code %d' % rc)
    return out


def decode_jpeg_coefficients(data: bytes, component: int) -> np.ndarray:
    """Quantised DCT coefficients of one component: (blocks_h, blocks_w, 64) int16, natural order,
    over the MCU-padded block grid."""
    data = bytes(data)
    (bw, bh) = (C.c_int(), C.c_int())
    rc = _jpeg().jo_decode_coefficients(data, len(data), component, None, 0, bw, bh)
    if rc:
        raise JpegError('jpeg oracle: code %d' % rc)
    out = np.empty((bh.value, bw.value, 64), np.int16)
    rc = _jpeg().jo_decode_coefficients(data, len(data), component, out.ctypes.data, out.size, bw, bh)
    if rc:
        raise JpegError('jpeg oracle: code %d' % rc)
    return out


# ---------------------------------------------------------------------------------------------
# oracle/_ref: the UNMODIFIED reference package, vendored at build time (git-ignored; it travels
# to the GPU box with the gpurun snapshot) and run one process per core as the timed CPU arm
# (ImageFile(name, params, bgr_image=crop) -> get_meter_value, _reading.py:19-115).
# ---------------------------------------------------------------------------------------------
REF_SOURCE = os.environ.get('MELF_REFERENCE_ROOT', '/root/reference')
REF_DIR = os.path.join(_HERE, '_ref')


def build_ref(force: bool = False) -> Optional[str]:
    """Copy ``/root/reference/meterelf`` (Python sources, nothing to compile) to ``oracle/_ref/meterelf``
    when the reference tree is present; returns the directory that holds the package, or None when
    there is neither a reference tree nor an earlier copy."""
    import shutil
    src = os.path.join(REF_SOURCE, 'meterelf')
    dst = os.path.join(REF_DIR, 'meterelf')
    if os.path.isdir(src) and (force or not os.path.isdir(dst)):
        if os.path.isdir(dst):
            shutil.rmtree(dst)
        os.makedirs(REF_DIR, exist_ok=True)
        shutil.copytree(src, dst, ignore=shutil.ignore_patterns('__pycache__', '*.pyc'))
        # the reference's own test file, run unchanged against the drop-in by
        # tests/test_dropin_reference_suite.py (the sample files come from the golden fixtures)
        os.makedirs(os.path.join(REF_DIR, 'tests'), exist_ok=True)
        shutil.copy(os.path.join(REF_SOURCE, 'tests', 'test_meterelf.py'), os.path.join(REF_DIR, 'tests'))
    return REF_DIR if os.path.isdir(dst) else None


def ref_available() -> bool:
    return os.path.isdir(os.path.join(REF_DIR, 'meterelf'))


class RefPool:
    """``workers`` processes of the unmodified reference (``oracle/ref_worker.py``), each reading a
    contiguous share of ``crops`` (n, h, w, 3 u8, already cut to meter_rect) with ``params_file``.
    ``run()`` starts all workers together and returns (wall seconds, records or None)."""

    def __init__(self, params_file: str, crops: np.ndarray, workers: int,
                 threshold: Optional[float] = None) -> None:
        import json
        import sys
        import tempfile
        if not ref_available():
            raise RuntimeError('oracle/_ref is absent: run __graft_entry__.build() where /root/reference exists')
        self._json = json
        self.n = len(crops)
        self.workers = max(1, min(workers, self.n))
        shm = '/dev/shm' if os.path.isdir('/dev/shm') and os.access('/dev/shm', os.W_OK) else None
        self._tmp = tempfile.TemporaryDirectory(dir=shm)
        path = os.path.join(self._tmp.name, 'crops.npy')
        np.save(path, np.ascontiguousarray(crops))
        env = dict(os.environ, OMP_NUM_THREADS='1', OPENBLAS_NUM_THREADS='1', MKL_NUM_THREADS='1',
                   CUDA_VISIBLE_DEVICES='')
        self.procs = []
        for k in range(self.workers):
            proc = subprocess.Popen([sys.executable, os.path.join(_HERE, 'ref_worker.py')],
                                    stdin=subprocess.PIPE, stdout=subprocess.PIPE, env=env, text=True)
            (lo, hi) = ((k * self.n) // self.workers, ((k + 1) * self.n) // self.workers)
            self._send(proc, {'cmd': 'load', 'root': REF_DIR, 'params': params_file, 'frames': path,
                              'lo': lo, 'hi': hi, 'threshold': threshold})
            self.procs.append(proc)
        for proc in self.procs:
            assert 'ready' in self._recv(proc)

    def _send(self, proc, msg) -> None:
        proc.stdin.write(self._json.dumps(msg) + '\n')
        proc.stdin.flush()

    def _recv(self, proc) -> Dict[str, Any]:
        line = proc.stdout.readline()
        if not line:
            raise RuntimeError('reference worker died (exit code %s)' % proc.poll())
        return self._json.loads(line)

    def run(self, want_results: bool = True):
        import time
        t0 = time.perf_counter()
        for proc in self.procs:
            self._send(proc, {'cmd': 'run', 'results': want_results})
        replies = [self._recv(proc) for proc in self.procs]
        wall = time.perf_counter() - t0
        records = None
        if want_results:
            records = [r for rep in replies for r in rep['results']]
            assert [r['i'] for r in records] == list(range(self.n))
        return (wall, records)

    def close(self) -> None:
        for proc in self.procs:
            try:
                self._send(proc, {'cmd': 'quit'})
                proc.stdin.close()
            except Exception:
                pass
        for proc in self.procs:
            try:
                proc.wait(timeout=10)
            except Exception:
                proc.kill()
        self.procs = []
        self._tmp.cleanup()

    def __enter__(self):
        return self

    def __exit__(self, *exc) -> None:
        self.close()


def ref_mismatches(ref_records: Sequence[Dict[str, Any]], got: np.ndarray, n_dials: int) -> int:
    """How many of ``got`` (result records of the GPU path or of the port) differ from what the
    reference workers returned for the same frames: status, failing / unreadable dials, float64 dial
    positions and value bit for bit."""
    bad = 0
    for (r, g) in zip(ref_records, got):
        same = int(g['status']) == r['status']
        if same and r['status'] == 0:
            want = [float.fromhex(v) for v in r['dial_pos']]
            same = [float(v) for v in g['dial_pos'][:n_dials]] == want and (
                r['value'] is None or float(g['value']) == float.fromhex(r['value']))
        elif same and r['status'] == 2:
            same = int(g['failed_dial']) == r['failed_dial']
        elif same and r['status'] == 3:
            same = int(g['unreadable_mask']) == r['unreadable_mask']
        bad += 0 if same else 1
    return bad
