"""ctypes binding of the C ABI (``include/meterelf_b200.h``) + the batching engine.

This is the only module that talks to ``libmeterelf_b200.so``.  There is NO CPU
fallback: if the CUDA library is missing or no GPU is present, every compute entry
point raises ``EngineUnavailableError``.
"""
import ctypes as C
import os
import threading
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from ._types import DialData, Image

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get(
    'MELF_LIB', os.path.join(_HERE, '_lib', 'libmeterelf_b200.so'))

MAX_DIALS = 8
ABI_VERSION = 1

ST_OK = 0
ST_DIALS_NOT_FOUND = 1
ST_NO_CONTOURS = 2
ST_ANGLE_UNDETERMINED = 3

ERR_NAMES = {1: 'MELF_ERR_INVALID', 2: 'MELF_ERR_CUDA', 3: 'MELF_ERR_UNSUPPORTED',
             4: 'MELF_ERR_TOO_SMALL'}
ERR_INVALID = 1
ERR_UNSUPPORTED = 3
ERR_TOO_SMALL = 4


class EngineUnavailableError(RuntimeError):
    """The CUDA extension (or a GPU) is missing; there is deliberately no fallback."""


class EngineError(RuntimeError):
    def __init__(self, code: int, text: str) -> None:
        super().__init__('{}: {}'.format(ERR_NAMES.get(code, code), text))
        self.code = code


class CDial(C.Structure):
    _fields_ = [('center_x', C.c_double), ('center_y', C.c_double),
                ('angle_of_zero', C.c_double),
                ('range_h', C.c_int32), ('range_l', C.c_int32), ('range_s', C.c_int32),
                ('negative_momentum', C.c_int32),
                ('mask', C.c_void_p), ('circle_mask', C.c_void_p)]


class CParams(C.Structure):
    _fields_ = [('struct_size', C.c_uint32), ('n_dials', C.c_int32),
                ('tmpl_w', C.c_int32), ('tmpl_h', C.c_int32), ('tmpl', C.c_void_p),
                ('rect_x0', C.c_int32), ('rect_y0', C.c_int32),
                ('rect_x1', C.c_int32), ('rect_y1', C.c_int32),
                ('hue_shift', C.c_int32), ('s_tail', C.c_int32),
                ('match_threshold', C.c_double), ('value_order', C.c_int32 * 4),
                ('dials', CDial * MAX_DIALS)]


class CTaps(C.Structure):
    _fields_ = [('d_mask_orig', C.c_void_p), ('d_mask_de', C.c_void_p),
                ('d_needle_mask', C.c_void_p), ('d_dial_color', C.c_void_p),
                ('d_counts', C.c_void_p), ('d_momentum', C.c_void_p)]


RESULT_DTYPE = np.dtype([
    ('status', np.int32), ('match_x', np.int32), ('match_y', np.int32),
    ('max_val', np.float32), ('failed_dial', np.int32), ('unreadable_mask', np.uint32),
    ('dial_pos', np.float64, (MAX_DIALS,)), ('value', np.float64)], align=True)

_lib: Optional[C.CDLL] = None
_lib_lock = threading.Lock()

_SYMBOLS = (
    'melf_create', 'melf_destroy', 'melf_last_error', 'melf_abi_version',
    'melf_dial_window', 'melf_read_batch_async', 'melf_read_batch', 'melf_convert_hls',
    'melf_match_template', 'melf_launch_count', 'melf_profile_enable', 'melf_profile_read',
    'melf_average_translated', 'melf_decode_jpeg', 'melf_jpeg_info', 'melf_read_jpeg_batch',
    'melf_in_range', 'melf_file_sizes', 'melf_read_files', 'melf_jpeg_probe_batch', 'melf_last_locate',
    'melf_match_template_noalloc')

NOALLOC_MAX_FRAMES = 8192           # MELF_NOALLOC_MAX_FRAMES

PROFILE_KINDS = ('match_prep', 'match_tc', 'dials', 'match_dp4a', 'jpeg_upload', 'jpeg_destuff',
                 'jpeg_huffman', 'jpeg_idct', 'jpeg_color')


def load_library() -> C.CDLL:
    """dlopen the CUDA library (no GPU needed for this step)."""
    global _lib
    with _lib_lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise EngineUnavailableError(
                'CUDA extension not built: {} (run `python -c "import __graft_entry__ as g; '
                'g.build()"`); there is no CPU fallback'.format(LIB_PATH))
        try:
            lib = C.CDLL(LIB_PATH)
        except OSError as error:
            raise EngineUnavailableError(
                'cannot load {}: {}'.format(LIB_PATH, error)) from error
        for name in _SYMBOLS:
            getattr(lib, name)      # AttributeError if the ABI is incomplete
        lib.melf_abi_version.restype = C.c_int
        if lib.melf_abi_version() != ABI_VERSION:
            raise EngineUnavailableError('ABI version mismatch in ' + LIB_PATH)
        lib.melf_last_error.restype = C.c_char_p
        lib.melf_last_error.argtypes = [C.c_void_p]
        lib.melf_launch_count.restype = C.c_uint64
        lib.melf_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(CParams)]
        lib.melf_destroy.argtypes = [C.c_void_p]
        lib.melf_destroy.restype = None
        lib.melf_dial_window.argtypes = [
            C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        lib.melf_read_batch_async.argtypes = [
            C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
            C.c_void_p, C.POINTER(CTaps), C.c_void_p]
        lib.melf_read_batch.argtypes = [
            C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
            C.c_void_p]
        lib.melf_convert_hls.argtypes = [
            C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
            C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
            C.c_void_p]
        lib.melf_match_template.argtypes = [
            C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t,
            C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.melf_match_template_noalloc.argtypes = [
            C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t,
            C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.melf_average_translated.argtypes = [
            C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_void_p,
            C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.melf_decode_jpeg.argtypes = [
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
            C.c_size_t, C.c_void_p]
        lib.melf_jpeg_info.argtypes = [C.c_void_p, C.c_size_t] + [C.POINTER(C.c_int)] * 5
        lib.melf_read_jpeg_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        lib.melf_in_range.argtypes = [
            C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_void_p,
            C.c_void_p, C.c_void_p, C.c_void_p]
        lib.melf_file_sizes.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_void_p]
        lib.melf_read_files.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.melf_jpeg_probe_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        lib.melf_last_locate.restype = C.c_char_p
        lib.melf_last_locate.argtypes = [C.c_void_p]
        lib.melf_profile_enable.argtypes = [C.c_void_p, C.c_int]
        lib.melf_profile_read.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
        _lib = lib
        return lib


def build_cparams(template: Image, meter_rect: Tuple[int, int, int, int], hue_shift: int,
                  match_threshold: float, dials: Sequence[Dict[str, Any]],
                  s_tail: bool = True) -> Tuple[CParams, List[Any]]:
    """Plain data -> ``melf_params``.  ``dials``: YAML-ordered dicts with keys name,
    center, color_range (h, l, s), negative_momentum, angle_of_zero, mask, circle_mask.
    Returns the struct and the list of arrays that must stay alive with it."""
    if len(dials) > MAX_DIALS:     # no dials = a locate-only context (match_template with any template)
        raise ValueError('at most {} dials are supported'.format(MAX_DIALS))
    keep: List[Any] = []
    tmpl = np.ascontiguousarray(template, dtype=np.uint8)
    if tmpl.ndim != 2:
        raise ValueError('dials template must be a single-channel image')
    keep.append(tmpl)
    p = CParams()
    p.struct_size = C.sizeof(CParams)
    p.n_dials = len(dials)
    (p.tmpl_h, p.tmpl_w) = tmpl.shape
    p.tmpl = tmpl.ctypes.data
    (p.rect_x0, p.rect_y0, p.rect_x1, p.rect_y1) = (int(v) for v in meter_rect)
    p.hue_shift = int(hue_shift)
    p.s_tail = 1 if s_tail else 0
    p.match_threshold = float(match_threshold)
    names = [d['name'] for d in dials]
    by_name = sorted(range(len(dials)), key=lambda i: names[i])   # _reading.py:171
    for (k, i) in enumerate(by_name[:4]):
        p.value_order[k] = i
    for (i, d) in enumerate(dials):
        mask = np.ascontiguousarray(d['mask'], dtype=np.uint8)
        ring = np.ascontiguousarray(d['circle_mask'], dtype=np.uint8)
        if mask.shape != tmpl.shape or ring.shape != tmpl.shape:
            raise ValueError('dial masks must have the dials template shape')
        keep += [mask, ring]
        q = p.dials[i]
        (q.center_x, q.center_y) = (float(d['center'][0]), float(d['center'][1]))
        q.angle_of_zero = float(d['angle_of_zero'])
        (q.range_h, q.range_l, q.range_s) = (int(v) for v in d['color_range'])
        q.negative_momentum = 1 if d['negative_momentum'] else 0
        q.mask = mask.ctypes.data
        q.circle_mask = ring.ctypes.data
    return (p, keep)


def dials_from_params(params: Any, dial_data: Dict[str, DialData]) -> List[Dict[str, Any]]:
    return [dict(
        name=name, center=dd.center,
        color_range=tuple(int(c) for c in params.dial_color_range[name]),
        negative_momentum=name in params.negative_momentum_dials,
        angle_of_zero=params.needle_angles_of_zero[name],
        mask=dd.mask, circle_mask=dd.circle_mask) for (name, dd) in dial_data.items()]


class Engine:
    """One ``melf_ctx``: a params set bound to one GPU."""

    def __init__(self, cparams: CParams, keep: List[Any], names: Sequence[str],
                 device: int = 0) -> None:
        self.lib = load_library()
        self.cparams = cparams
        self._keep = keep
        self.names = list(names)
        self.n_dials = cparams.n_dials
        self.device = device
        self._call_lock = threading.Lock()
        self._ctx = C.c_void_p()
        rc = self.lib.melf_create(C.byref(self._ctx), device, C.byref(cparams))
        if rc:
            text = (self.lib.melf_last_error(None) or b'').decode()
            if rc == 2:
                raise EngineUnavailableError(
                    'CUDA device {} unavailable: {} (no CPU fallback)'.format(device, text))
            raise EngineError(rc, text)

    def close(self) -> None:
        if self._ctx:
            self.lib.melf_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self) -> None:
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int) -> None:
        if rc:
            raise EngineError(rc, (self.lib.melf_last_error(self._ctx) or b'').decode())

    def profile_enable(self, on: bool) -> None:
        self._check(self.lib.melf_profile_enable(self._ctx, 1 if on else 0))

    def profile_read(self) -> Dict[str, Tuple[float, int]]:
        """kernel kind -> (milliseconds, launches) since the last read (synchronises)."""
        ms = (C.c_float * len(PROFILE_KINDS))()
        cnt = (C.c_int32 * len(PROFILE_KINDS))()
        self._check(self.lib.melf_profile_read(self._ctx, ms, cnt))
        return {k: (float(ms[i]), int(cnt[i])) for (i, k) in enumerate(PROFILE_KINDS)}

    def last_locate(self) -> str:
        """Which locate kernel the last call took (size dispatch inside the library)."""
        return (self.lib.melf_last_locate(self._ctx) or b'').decode()

    def dial_window(self, dial: int) -> Tuple[int, int]:
        x = C.c_int32()
        y = C.c_int32()
        self._check(self.lib.melf_dial_window(self._ctx, dial, C.byref(x), C.byref(y)))
        return (x.value, y.value)

    # -- host frames in, host results out: what the Python API uses ------------------
    def read_frames(self, frames: np.ndarray) -> np.ndarray:
        """frames: (n, h, w, 3) u8 host array (rows contiguous); returns RESULT_DTYPE[n]."""
        if frames.dtype != np.uint8 or frames.ndim != 4 or frames.shape[3] != 3:
            raise ValueError('frames must be (n, h, w, 3) uint8')
        if frames.strides[3] != 1 or frames.strides[2] != 3:
            frames = np.ascontiguousarray(frames)
        (n, h, w) = frames.shape[:3]
        out = np.zeros((n,), RESULT_DTYPE)
        if n:
            with self._call_lock:           # one batch at a time per context (its staging and slots)
                self._check(self.lib.melf_read_batch(
                    self._ctx, frames.ctypes.data, n, h, w, frames.strides[1],
                    frames.strides[0], out.ctypes.data))
        return out

    # -- JPEG streams in (host), results out (host): SURVEY 8f-1 -------------------------
    def read_jpeg_streams(self, streams: Sequence[bytes]) -> np.ndarray:
        """``get_meter_values`` over JPEG files already read into memory; RESULT_DTYPE[n]."""
        (blob, offsets) = pack_streams(streams)
        return self.read_jpeg_blob(blob, offsets)

    def read_jpeg_blob(self, blob: np.ndarray, offsets: np.ndarray) -> np.ndarray:
        """The same for streams that are already packed: stream i = blob[offsets[i]:offsets[i+1]]."""
        n = len(offsets) - 1
        out = np.zeros((n,), RESULT_DTYPE)
        if n:
            with self._call_lock:
                self._check(self.lib.melf_read_jpeg_batch(
                    self._ctx, blob.ctypes.data, offsets.ctypes.data, n, out.ctypes.data))
        return out

    def decode_jpeg_device(self, blob: np.ndarray, offsets: np.ndarray, h: int, w: int, d_bgr: int,
                           row_stride: int, frame_stride: int, stream: int = 0) -> None:
        """``cv2.imread`` of len(offsets)-1 JPEG streams into device BGR frames."""
        self._check(self.lib.melf_decode_jpeg(
            self._ctx, blob.ctypes.data, offsets.ctypes.data, len(offsets) - 1, h, w, d_bgr,
            row_stride, frame_stride, stream))

    # -- device frames in, device results out (raw pointers; async on `stream`) -------
    def read_device(self, d_bgr: int, n: int, h: int, w: int, row_stride: int,
                    frame_stride: int, d_out: int, stream: int = 0,
                    taps: Optional[CTaps] = None) -> None:
        self._check(self.lib.melf_read_batch_async(
            self._ctx, d_bgr, n, h, w, row_stride, frame_stride, d_out,
            C.byref(taps) if taps is not None else None, stream))

    def match_template_device(self, d_img: int, channels: int, n: int, h: int, w: int,
                              row_stride: int, frame_stride: int, use_rect: bool,
                              d_max_xy: int, d_max_val: int, d_result: int = 0,
                              stream: int = 0) -> None:
        self._check(self.lib.melf_match_template(
            self._ctx, d_img, channels, n, h, w, row_stride, frame_stride,
            1 if use_rect else 0, d_max_xy, d_max_val, d_result or None, stream))

    def match_template_noalloc_device(self, d_img: int, channels: int, n: int, h: int, w: int,
                                      row_stride: int, frame_stride: int, use_rect: bool,
                                      d_max_xy: int, d_max_val: int, stream: int = 0) -> None:
        """``melf_match_template_noalloc``: the same arg-max on the library's static scratch (no cudaMalloc;
        at most ``NOALLOC_MAX_FRAMES`` frames per call; calls that overlap on the GPU race)."""
        self._check(self.lib.melf_match_template_noalloc(
            self._ctx, d_img, channels, n, h, w, row_stride, frame_stride,
            1 if use_rect else 0, d_max_xy, d_max_val, stream))


def convert_hls_device(device: int, d_bgr: int, n: int, h: int, w: int, row_stride: int,
                       frame_stride: int, hue_shift: int, s_tail: bool,
                       d_hls: int = 0, d_light: int = 0,
                       lo: Optional[Sequence[int]] = None, hi: Optional[Sequence[int]] = None,
                       d_mask: int = 0, stream: int = 0) -> None:
    lib = load_library()
    lo_b = (C.c_uint8 * 3)(*(lo or (0, 0, 0)))
    hi_b = (C.c_uint8 * 3)(*(hi or (255, 255, 255)))
    rc = lib.melf_convert_hls(device, d_bgr, n, h, w, row_stride, frame_stride, hue_shift,
                              1 if s_tail else 0, d_hls or None, d_light or None,
                              lo_b, hi_b, d_mask or None, stream or None)
    if rc:
        raise EngineError(rc, (lib.melf_last_error(None) or b'').decode())


def in_range_device(device: int, d_img: int, n: int, h: int, w: int, row_stride: int, frame_stride: int,
                    lo: Sequence[int], hi: Sequence[int], d_mask: int, stream: int = 0) -> None:
    """``melf_in_range``: cv2.inRange of 3-channel u8 frames on the device."""
    lib = load_library()
    lo_b = (C.c_uint8 * 3)(*(int(v) for v in lo))
    hi_b = (C.c_uint8 * 3)(*(int(v) for v in hi))
    rc = lib.melf_in_range(device, d_img, n, h, w, row_stride, frame_stride, lo_b, hi_b, d_mask,
                           stream or None)
    if rc:
        raise EngineError(rc, (lib.melf_last_error(None) or b'').decode())


def pack_streams(streams: Sequence[bytes]) -> Tuple[np.ndarray, np.ndarray]:
    """Concatenates byte strings into (blob u8, offsets u64[n+1]) as the JPEG entry points take."""
    sizes = np.array([len(s) for s in streams], np.uint64)
    offsets = np.zeros(len(streams) + 1, np.uint64)
    np.cumsum(sizes, out=offsets[1:])
    blob = np.frombuffer(b''.join(bytes(s) for s in streams) or b'\0', np.uint8)
    return (blob, offsets)


class _Blob(np.ndarray):
    """ndarray view of a pooled page-locked buffer (``_owner``: the torch tensor that holds the memory)."""
    _owner: Any = None


_blob_lock = threading.Lock()
_blob_free: List[Any] = []                  # page-locked torch tensors waiting for the next batch


def _blob_buffer(nbytes: int) -> np.ndarray:
    """``nbytes`` of host memory for a batch of file bytes.  With a GPU it is page-locked and comes from a
    small pool (``release_blob`` hands it back): the ingest path then copies it to the device as it lies --
    pageable memory is first copied into a pinned staging area, ~40 ms per 300 MB batch -- and fresh pages are
    not faulted in for every batch."""
    try:
        import torch
        pinned = torch.cuda.is_available()
    except ImportError:
        pinned = False
    if not pinned:
        return np.empty((nbytes,), np.uint8)
    buf = None
    with _blob_lock:
        fits = [k for (k, t) in enumerate(_blob_free) if t.numel() >= nbytes]
        if fits:
            buf = _blob_free.pop(min(fits, key=lambda k: _blob_free[k].numel()))
    if buf is None:
        buf = torch.empty((max(nbytes + nbytes // 4, 1 << 20),), dtype=torch.uint8, pin_memory=True)
    arr = buf.numpy()[:nbytes].view(_Blob)
    arr._owner = buf
    return arr


def release_blob(blob: np.ndarray) -> None:
    """Hands the buffer of a ``read_files`` blob back to the pool; the blob must not be used afterwards."""
    owner = getattr(blob, '_owner', None)
    if owner is None:
        return
    blob._owner = None
    with _blob_lock:
        if len(_blob_free) < 4:
            _blob_free.append(owner)


def read_files(paths: Sequence[str]) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """The bytes of ``paths`` in one blob: (blob u8, offsets u64[n+1], ok bool[n]); a file that cannot be
    opened or read in full has ok False and an empty slice (``cv2.imread`` would return None for it)."""
    lib = load_library()
    n = len(paths)
    c_paths = (C.c_char_p * n)(*(os.fsencode(p) for p in paths))
    sizes = np.zeros((n,), np.int64)
    lib.melf_file_sizes(c_paths, n, sizes.ctypes.data)
    offsets = np.zeros(n + 1, np.uint64)
    np.cumsum(np.maximum(sizes, 0), out=offsets[1:])
    blob = _blob_buffer(int(offsets[-1]) + 16)
    ok = np.zeros((n,), np.int32)
    lib.melf_read_files(c_paths, n, blob.ctypes.data, offsets.ctypes.data, ok.ctypes.data)
    return (blob, offsets, (ok != 0) & (sizes >= 0))


JPEG_INFO_FIELDS = ('h', 'w', 'components', 'h_samp', 'v_samp', 'orientation', 'has_eoi', 'code')


def jpeg_probe_batch(blob: np.ndarray, offsets: np.ndarray) -> np.ndarray:
    """Header probe of every packed stream (no GPU): int32 (n, 8), columns ``JPEG_INFO_FIELDS``."""
    n = len(offsets) - 1
    info = np.zeros((n, 8), np.int32)
    if n:
        load_library().melf_jpeg_probe_batch(blob.ctypes.data, offsets.ctypes.data, n, info.ctypes.data)
    return info


def jpeg_info(data: bytes) -> Dict[str, int]:
    """Header probe of one JPEG stream (no GPU needed)."""
    lib = load_library()
    data = bytes(data)
    vals = [C.c_int() for _ in range(5)]
    rc = lib.melf_jpeg_info(data, len(data), *[C.byref(v) for v in vals])
    if rc:
        raise EngineError(rc, (lib.melf_last_error(None) or b'').decode())
    return dict(zip(('h', 'w', 'components', 'h_samp', 'v_samp'), (v.value for v in vals)))


def average_translated_device(device: int, d_bgr: int, n: int, h: int, w: int, row_stride: int,
                              frame_stride: int, d_match_xy: int, anchor: Sequence[int],
                              d_out: int, stream: int = 0) -> None:
    """``melf_average_translated``: the average meter image of the offline calibration."""
    lib = load_library()
    rc = lib.melf_average_translated(device, d_bgr, n, h, w, row_stride, frame_stride, d_match_xy,
                                     int(anchor[0]), int(anchor[1]), d_out, stream or None)
    if rc:
        raise EngineError(rc, 'melf_average_translated failed')


def launch_count() -> int:
    return int(load_library().melf_launch_count())
