"""Pixel primitives (reference: ``_utils.py:14-119``) as GPU operators.

Each function keeps the reference's name and argument meaning and round-trips numpy
arrays through the C ABI (``melf_convert_hls``, ``melf_match_template``); torch is used
only to own the device buffers.  For bulk work use the batched path in ``_reading``.
"""
import math
from typing import Optional, Tuple

import numpy

from . import _engine
from ._colors import HlsColor
from ._types import FloatPoint, Image, Point, Rect, TemplateMatchResult


def float_point_to_int(point: FloatPoint) -> Point:
    return (int(round(point[0])), int(round(point[1])))


def get_angle_by_vector(vector: FloatPoint) -> Optional[float]:
    """Direction of a vector as a fraction of a turn: 0 = up, clockwise, image y down.

    >>> [get_angle_by_vector(v) for v in [(0, -1), (1, 0), (0, 1), (-1, 0), (0, 0)]]
    [0.0, 0.25, 0.5, 0.75, None]
    """
    (x, y) = vector
    if y == 0:
        return 0.25 if x > 0 else 0.75 if x < 0 else None
    turns = math.atan(x / y) / (2 * math.pi)
    return (-turns + (0.5 if y > 0 else 0.0)) % 1.0


def crop_rect(img: Image, rect: Rect) -> Image:
    ((x0, y0), (x1, y1)) = rect
    return img[y0:y1, x0:x1]


def _torch():
    try:
        import torch
    except ImportError as error:     # pragma: no cover
        raise _engine.EngineUnavailableError('torch is needed for device buffers') from error
    if not torch.cuda.is_available():
        raise _engine.EngineUnavailableError('no CUDA device (there is no CPU fallback)')
    return torch


def _upload(image: Image):
    torch = _torch()
    host = numpy.ascontiguousarray(image)
    return torch.from_numpy(host).cuda()


def convert_to_hls(image: Image, hue_shift: int = 0,
                   color_range: Optional[Tuple[HlsColor, HlsColor]] = None):
    """BGR -> HLS_FULL plus wrapping hue shift.  With ``color_range=(lo, hi)`` the fused
    inRange mask is returned as well: ``(hls, mask)``."""
    torch = _torch()
    d_in = _upload(image)
    (h, w) = image.shape[:2]
    d_hls = torch.empty((h, w, 3), dtype=torch.uint8, device='cuda')
    d_mask = torch.empty((h, w), dtype=torch.uint8, device='cuda') if color_range else None
    (lo, hi) = color_range if color_range else (None, None)
    _engine.convert_hls_device(
        torch.cuda.current_device(), d_in.data_ptr(), 1, h, w, w * 3, h * w * 3,
        hue_shift, True, d_hls=d_hls.data_ptr(),
        lo=[int(c) for c in lo] if color_range else None,
        hi=[int(c) for c in hi] if color_range else None,
        d_mask=d_mask.data_ptr() if color_range else 0,
        stream=torch.cuda.current_stream().cuda_stream)
    hls = d_hls.cpu().numpy()
    return (hls, d_mask.cpu().numpy()) if color_range else hls


def get_mask_by_color(hls_image: Image, color: HlsColor, color_range: HlsColor) -> Image:
    """``cv2.inRange`` of an HLS image around ``color`` (``_utils.py:113-119``): 255 where every
    channel lies within ``color +- color_range`` (clamped to 0..255, no hue wrap), else 0."""
    torch = _torch()
    (lo, hi) = color.get_range(color_range)
    d_in = _upload(hls_image)
    (h, w) = hls_image.shape[:2]
    d_mask = torch.empty((h, w), dtype=torch.uint8, device='cuda')
    _engine.in_range_device(torch.cuda.current_device(), d_in.data_ptr(), 1, h, w, w * 3, h * w * 3,
                            [int(c) for c in lo], [int(c) for c in hi], d_mask.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
    return d_mask.cpu().numpy()


def get_mask_by_color_bgr(bgr_image: Image, hue_shift: int, color: HlsColor,
                          color_range: HlsColor) -> Image:
    """``get_mask_by_color(convert_to_hls(bgr, hue_shift), ...)`` in one fused pass over the BGR
    image (the HLS image is never written)."""
    return convert_to_hls(bgr_image, hue_shift, color.get_range(color_range))[1]


_template_engines: dict = {}


def _template_engine(template: Image, device: int) -> '_engine.Engine':
    """Locate-only ``melf_ctx`` (no dials) for an arbitrary template, cached by content."""
    tmpl = numpy.ascontiguousarray(template, dtype=numpy.uint8)
    key = (tmpl.shape, hash(tmpl.tobytes()), device)
    entry = _template_engines.get(key)
    if entry is None or not numpy.array_equal(entry[0], tmpl):
        (cparams, keep) = _engine.build_cparams(tmpl, (0, 0, 2 ** 30, 2 ** 30), 0, 0.0, [])
        entry = (tmpl, _engine.Engine(cparams, keep, [], device))
        if len(_template_engines) > 16:
            _template_engines.clear()
        _template_engines[key] = entry
    return entry[1]


def _match(engine, image: Image, template_shape) -> TemplateMatchResult:
    torch = _torch()
    channels = 3 if image.ndim == 3 else 1
    d_in = _upload(image)
    (h, w) = image.shape[:2]
    d_xy = torch.empty((2,), dtype=torch.int32, device='cuda')
    d_val = torch.empty((1,), dtype=torch.float32, device='cuda')
    engine.match_template_device(
        d_in.data_ptr(), channels, 1, h, w, w * channels, h * w * channels, False, d_xy.data_ptr(),
        d_val.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    (x, y) = (int(v) for v in d_xy.cpu())
    (th, tw) = template_shape[:2]
    return TemplateMatchResult(Rect((x, y), (x + tw, y + th)), float(d_val.cpu()[0]))


def match_template(img: Image, template: Image) -> TemplateMatchResult:
    """``cv2.matchTemplate(img, template, TM_CCOEFF)`` + ``cv2.minMaxLoc`` on a single-channel
    8-bit image (``_utils.py:91-97``): rect of the best match and its score.  Exact integer
    arithmetic, the first maximum in raster order wins (DESIGN.md section 2 for how that relates to
    cv2's float32 DFT on near-ties)."""
    if img.ndim != 2 or template.ndim != 2:
        raise ValueError('match_template takes single-channel images')
    if img.shape[0] < template.shape[0] or img.shape[1] < template.shape[1]:
        raise ValueError('image smaller than the template')     # cv2.matchTemplate raises cv2.error
    torch = _torch()
    return _match(_template_engine(template, torch.cuda.current_device()), img, template.shape)


def match_template_bgr(bgr_image: Image, template: Image, params) -> TemplateMatchResult:
    """``match_template`` on the lightness plane of a BGR image, fused (the plane is never written)."""
    from ._reading import get_engine
    torch = _torch()
    return _match(get_engine(params, cropped_input=True, device=torch.cuda.current_device()),
                  bgr_image, template.shape)
