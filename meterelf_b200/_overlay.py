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
4x pictures.  The pictures are the reference's, pixel for pixel (``tests/golden/
make_golden_overlay.py`` captures what the unmodified reference hands to ``cv2.imshow``): the
background is the HLS dials rect converted back to BGR like ``cv2.cvtColor(HLS2BGR_FULL)`` does
(``hls_to_bgr_full``; 4 of the 2^24 HLS colours come out one level off in one channel, float32
near-ties of OpenCV's vectorised code), the 4x / 2x scaling is ``cv2.resize``'s fixed-point bilinear
filter (``resize_linear``).  The PNG encoder is ``zlib`` (no OpenCV needed).
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


def hls_to_bgr_full(hls: numpy.ndarray) -> numpy.ndarray:
    """``cv2.cvtColor(hls, cv2.COLOR_HLS2BGR_FULL)`` of an 8-bit image (``convert_to_bgr``,
    ``_utils.py:105-110``), in the float32 operation order of OpenCV's vectorised ``HLS2RGB_f``:
    hue range 255 for the inverse conversion (the forward one uses 256), p = l*s or s - l*s, the
    four candidates l +- p and l +- (p - 2 f p), picked by hue sector; rint(v * 255) at the end."""
    f32 = numpy.float32
    hue = hls[..., 0].astype(f32)
    light = (hls[..., 1].astype(f32) * f32(1.0 / 255.0)).astype(f32)
    sat = (hls[..., 2].astype(f32) * f32(1.0 / 255.0)).astype(f32)
    ls = (light * sat).astype(f32)
    p = numpy.where(light <= f32(0.5), ls, (sat - ls).astype(f32))
    raw = (hue * (f32(6.0) / f32(255.0))).astype(f32)
    whole = numpy.trunc(raw).astype(f32)
    frac = (raw - whole).astype(f32)
    sector = (whole - (f32(6.0) * numpy.trunc((raw * f32(1.0 / 6.0)).astype(f32)).astype(f32)).astype(f32)).astype(f32)
    twice = (frac + frac).astype(f32)
    tab0 = (light + p).astype(f32)
    tab1 = (light - p).astype(f32)
    tab2 = (light + (p - (twice * p).astype(f32)).astype(f32)).astype(f32)
    tab3 = ((light - p).astype(f32) + (twice * p).astype(f32)).astype(f32)
    blue = numpy.where(sector < 2, tab1, numpy.where(sector <= 2, tab3, numpy.where(sector <= 4, tab0, tab2)))
    green = numpy.where(sector < 1, tab3, numpy.where(sector <= 2, tab0, numpy.where(sector < 4, tab2, tab1)))
    red = numpy.where(sector < 1, tab0, numpy.where(sector < 2, tab2, numpy.where(
        sector < 4, tab1, numpy.where(sector <= 4, tab3, tab0))))
    bgr = numpy.rint((numpy.stack([blue, green, red], axis=-1) * f32(255.0)).astype(f32))
    return numpy.clip(bgr, 0, 255).astype(numpy.uint8)


def resize_linear(img: numpy.ndarray, scale: int) -> numpy.ndarray:
    """``cv2.resize(img, (w * scale, h * scale))`` (INTER_LINEAR) of an 8-bit image: sample positions
    (d + 0.5) / scale - 0.5, 11-bit fixed-point coefficients, horizontal pass first; at the left / right
    edge the fraction is dropped, at the top / bottom only the row index is clamped (so the truncating
    fixed-point sum of the two equal rows stays what OpenCV computes)."""
    assert scale > 0
    (h, w) = img.shape[:2]
    src = img.reshape(h, w, -1).astype(numpy.int64)
    pos = (numpy.arange(w * scale) + 0.5) / scale - 0.5
    x0 = numpy.floor(pos).astype(numpy.int64)
    fx = pos - x0
    fx = numpy.where((x0 < 0) | (x0 >= w - 1), 0.0, fx)
    x0 = numpy.clip(x0, 0, w - 1)
    x1 = numpy.minimum(x0 + 1, w - 1)
    (ax0, ax1) = (numpy.rint((1 - fx) * 2048).astype(numpy.int64), numpy.rint(fx * 2048).astype(numpy.int64))
    pos = (numpy.arange(h * scale) + 0.5) / scale - 0.5
    y0 = numpy.floor(pos).astype(numpy.int64)
    fy = pos - y0
    (ay0, ay1) = (numpy.rint((1 - fy) * 2048).astype(numpy.int64), numpy.rint(fy * 2048).astype(numpy.int64))
    (y0, y1) = (numpy.clip(y0, 0, h - 1), numpy.clip(y0 + 1, 0, h - 1))
    rows = src[:, x0] * ax0[None, :, None] + src[:, x1] * ax1[None, :, None]
    out = (((ay0[:, None, None] * (rows[y0] >> 4)) >> 16) + ((ay1[:, None, None] * (rows[y1] >> 4)) >> 16) + 2) >> 2
    return out.astype(numpy.uint8).reshape((h * scale, w * scale) + img.shape[2:])


def scale_image(img: numpy.ndarray, scale: int) -> numpy.ndarray:
    """``scale_image`` of the reference (``_utils.py:57-61``)."""
    return resize_linear(img, scale)


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
                hls: Optional[numpy.ndarray] = None,
                ) -> Tuple[Optional[numpy.ndarray], List[Tuple[str, numpy.ndarray]]]:
    """(debug image of the dials rect, [(dial name, 4x picture after that dial)]) for one frame;
    (None, []) when the dials were not found (the reference raises before any window opens)."""
    if int(record['status']) == _engine.ST_DIALS_NOT_FOUND:
        return (None, [])
    (th, tw) = params.dials_template_size                 # (rows, columns), like ndarray.shape
    (mx, my) = (int(record['match_x']), int(record['match_y']))
    if hls is None:
        from ._utils import convert_to_hls
        hls = convert_to_hls(numpy.ascontiguousarray(crop), params.hue_shift)
    # convert_to_bgr (_utils.py:105-110): the hue shift taken off again (wrapping u8), HLS -> BGR
    dials_hls = numpy.array(hls[my:my + th, mx:mx + tw], dtype=numpy.uint8)
    dials_hls[..., 0] -= numpy.uint8(params.hue_shift)
    debug = hls_to_bgr_full(dials_hls)
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
        if int(records[i]['status']) != _engine.ST_OK:
            continue                      # the reference raises before its last imshow (_reading.py:98-106)
        written.append(stem + '.png')
        with open(written[-1], 'wb') as fp:
            fp.write(encode_png(scale_image(debug, 2)))
    return written
