"""Debug overlays as image files (reference: ``_reading.py:43-78, 112-114, 143-144``).

The reference paints its intermediate results into a BGR copy of the dials rect and shows it
with ``cv2.imshow`` (one window per dial at 4x, one per file at 2x).  There is no display
here, so the same pictures are written as PNG files instead: with ``DEBUG`` set and
``MELF_DEBUG_DIR=<dir>``, ``get_meter_values`` leaves ``<dir>/<image name>.png`` (2x) and
``<dir>/<image name>.dial-<dial name>.png`` (4x, cumulative like the reference's windows).

What is painted comes from the GPU path's own taps (``melf_taps``: selected needle mask,
momentum vector, contour area), not from a second implementation: filled needle contour in
(255, 255, 0) when its area exceeds 100, the momentum marker (radius-4 circle, red) 24 px from
the dial centre, ring pixels of the needle in (0, 128, 128) and the ones that survive the
momentum filter in (0, 255, 255), the dial centre (black dot, magenta radius-6 circle) in the
4x pictures.  Differences to the reference's windows: the background is the decoded BGR
crop itself (the reference converts its HLS image back to BGR), scaling is pixel replication
(``cv2.resize`` interpolates), and the PNG encoder is ``zlib`` (no OpenCV needed).
"""
import math
import os
import struct
import zlib
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy

from . import _debug, _engine
from ._dial_data import draw_circle, get_dial_data
from ._params import Params as _Params
from ._utils import float_point_to_int, get_angle_by_vector

BGR_NEEDLE = (255, 255, 0)
BGR_MOMENTUM = (0, 0, 255)
BGR_RING = (0, 128, 128)
BGR_RING_KEPT = (0, 255, 255)
BGR_BLACK = (0, 0, 0)
BGR_MAGENTA = (255, 0, 255)


def overlay_dir() -> Optional[str]:
    """Where the pictures go, or None when the feature is off."""
    path = os.getenv('MELF_DEBUG_DIR', '')
    return path if (path and _debug.DEBUG) else None


def encode_png(bgr: numpy.ndarray) -> bytes:
    """8-bit RGB PNG (filter 0 on every row) of a (h, w, 3) BGR image."""
    (h, w) = bgr.shape[:2]
    rgb = numpy.ascontiguousarray(bgr[:, :, ::-1], dtype=numpy.uint8)
    raw = numpy.zeros((h, 1 + 3 * w), numpy.uint8)
    raw[:, 1:] = rgb.reshape(h, 3 * w)

    def chunk(kind: bytes, body: bytes) -> bytes:
        return (struct.pack('>I', len(body)) + kind + body +
                struct.pack('>I', zlib.crc32(kind + body) & 0xFFFFFFFF))

    return (b'\x89PNG\r\n\x1a\n' + chunk(b'IHDR', struct.pack('>IIBBBBB', w, h, 8, 2, 0, 0, 0)) +
            chunk(b'IDAT', zlib.compress(raw.tobytes(), 6)) + chunk(b'IEND', b''))


def scale_image(img: numpy.ndarray, scale: int) -> numpy.ndarray:
    assert scale > 0
    return numpy.repeat(numpy.repeat(img, scale, axis=0), scale, axis=1)


def window_to_mask(words: numpy.ndarray, win_x0: int, win_y0: int, shape: Tuple[int, int]) -> numpy.ndarray:
    """A 64x64 tap window (row r = 64-bit word, bit c = pixel (win_x0 + c, win_y0 + r)) as a
    boolean image of the dials rect."""
    (h, w) = shape
    out = numpy.zeros((h, w), bool)
    rows = numpy.asarray(words, numpy.uint64)
    bits = ((rows[:, None] >> numpy.arange(64, dtype=numpy.uint64)[None, :]) & numpy.uint64(1)).astype(bool)
    (y0, y1) = (max(win_y0, 0), min(win_y0 + 64, h))
    (x0, x1) = (max(win_x0, 0), min(win_x0 + 64, w))
    if y1 > y0 and x1 > x0:
        out[y0:y1, x0:x1] = bits[y0 - win_y0:y1 - win_y0, x0 - win_x0:x1 - win_x0]
    return out


def read_with_taps(engine: _engine.Engine, frames: numpy.ndarray) -> Tuple[numpy.ndarray, Dict[str, numpy.ndarray]]:
    """``melf_read_batch_async`` with every tap attached: (records, taps as host arrays)."""
    from ._utils import _torch
    torch = _torch()
    frames = numpy.ascontiguousarray(frames)
    (n, h, w) = frames.shape[:3]
    nd = engine.n_dials
    d_in = torch.from_numpy(frames).cuda()
    d_out = torch.zeros((n, _engine.RESULT_DTYPE.itemsize), dtype=torch.uint8, device='cuda')
    words = [torch.zeros((n, nd, 64), dtype=torch.int64, device='cuda') for _ in range(3)]
    color = torch.zeros((n, nd, 3), dtype=torch.int32, device='cuda')
    counts = torch.zeros((n, nd, 4), dtype=torch.int32, device='cuda')
    momentum = torch.zeros((n, nd, 3), dtype=torch.float64, device='cuda')
    taps = _engine.CTaps(words[0].data_ptr(), words[1].data_ptr(), words[2].data_ptr(),
                         color.data_ptr(), counts.data_ptr(), momentum.data_ptr())
    engine.read_device(d_in.data_ptr(), n, h, w, w * 3, h * w * 3, d_out.data_ptr(),
                       torch.cuda.current_stream().cuda_stream, taps)
    torch.cuda.synchronize()
    records = d_out.cpu().numpy().view(_engine.RESULT_DTYPE).reshape(n)
    return (records, {
        'needle_mask': words[2].cpu().numpy().view(numpy.uint64),
        'counts': counts.cpu().numpy(), 'momentum': momentum.cpu().numpy()})


def paint_frame(params: _Params, engine: _engine.Engine, crop: numpy.ndarray, record: Any,
                needle_words: numpy.ndarray, counts: numpy.ndarray, momentum: numpy.ndarray,
                ) -> Tuple[Optional[numpy.ndarray], List[Tuple[str, numpy.ndarray]]]:
    """(debug image of the dials rect, [(dial name, 4x picture after that dial)]) for one frame;
    (None, []) when the dials were not found (the reference raises before any window opens)."""
    if int(record['status']) == _engine.ST_DIALS_NOT_FOUND:
        return (None, [])
    (th, tw) = params.dials_template_size                 # (rows, columns), like ndarray.shape
    (mx, my) = (int(record['match_x']), int(record['match_y']))
    debug = numpy.array(crop[my:my + th, mx:mx + tw], dtype=numpy.uint8)
    (h, w) = debug.shape[:2]
    per_dial: List[Tuple[str, numpy.ndarray]] = []
    for (d, (name, dial)) in enumerate(get_dial_data(params).items()):
        if int(record['status']) == _engine.ST_NO_CONTOURS and int(record['failed_dial']) == d:
            break                                         # the reference raises here (_reading.py:137-138)
        (win_x0, win_y0) = engine.dial_window(d)
        needle = window_to_mask(needle_words[d], win_x0, win_y0, (h, w))
        if int(counts[d][2]) > 200:                       # 2 * contourArea > 2 * 100 (_reading.py:142-144)
            debug[needle] = BGR_NEEDLE
        (mom_x, mom_y, mom_angle) = (float(v) for v in momentum[d])
        sign = -1 if name in params.negative_momentum_dials else 1
        norm = math.sqrt(mom_x ** 2 + mom_y ** 2)
        (cx, cy) = dial.center
        if norm > 0:
            marker = float_point_to_int((cx + 24 * sign * mom_x / norm, cy + 24 * sign * mom_y / norm))
            draw_circle(debug, marker, 4, BGR_MOMENTUM)
        ring = needle & (numpy.asarray(dial.circle_mask)[:h, :w] != 0)
        for (y, x) in zip(*numpy.nonzero(ring)):          # raster order, like findNonZero
            debug[y, x] = BGR_RING
            angle = get_angle_by_vector((x - cx, y - cy))
            if angle is None or math.isnan(mom_angle):
                continue
            dist = abs(angle - mom_angle)
            if min(dist, abs(dist - 1)) < 0.25:
                debug[y, x] = BGR_RING_KEPT
        big = scale_image(debug, 4)
        centre = float_point_to_int((cx * 4, cy * 4))
        if 0 <= centre[0] < big.shape[1] and 0 <= centre[1] < big.shape[0]:
            big[centre[1], centre[0]] = BGR_BLACK
        draw_circle(big, centre, 6, BGR_MAGENTA)
        per_dial.append((name, big))
    return (debug, per_dial)


def write_overlays(params: _Params, engine: _engine.Engine, frames: numpy.ndarray,
                   filenames: Sequence[str], directory: str) -> List[str]:
    """Re-reads ``frames`` (already cut to ``meter_rect``) with the taps attached and writes the
    pictures; returns the paths written."""
    os.makedirs(directory, exist_ok=True)
    (records, taps) = read_with_taps(engine, frames)
    written: List[str] = []
    for (i, filename) in enumerate(filenames):
        (debug, per_dial) = paint_frame(params, engine, frames[i], records[i], taps['needle_mask'][i],
                                        taps['counts'][i], taps['momentum'][i])
        if debug is None:
            continue
        stem = os.path.join(directory, os.path.basename(filename))
        for (name, picture) in per_dial:
            written.append('{}.dial-{}.png'.format(stem, name))
            with open(written[-1], 'wb') as fp:
                fp.write(encode_png(picture))
        written.append(stem + '.png')
        with open(written[-1], 'wb') as fp:
            fp.write(encode_png(scale_image(debug, 2)))
    return written
