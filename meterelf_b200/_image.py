"""Frame I/O and dial location (reference: ``_image.py:12-81``).

Files that are baseline JPEGs the ingest kernels support (the pixels are
bit-identical to ``cv2.imread``) are handed to the GPU as bytes (``read_stream`` +
``gpu_ingest_geometry``); everything else -- PNG, progressive or EXIF-rotated JPEGs,
injected ``bgr_image`` arrays -- is decoded with ``cv2.imread`` exactly as in the
reference and uploaded as pixels.  ``MELF_INGEST=host`` forces the latter for every file.
"""
import os
from typing import Dict, Optional, Tuple

import numpy

from ._params import Params as _Params
from ._types import Image, Rect, TemplateMatchResult
from .exceptions import DialsNotFoundError, ImageLoadingError


def _imread(filename: str, grayscale: bool = False) -> Optional[Image]:
    import cv2
    return cv2.imread(filename, cv2.IMREAD_GRAYSCALE if grayscale else cv2.IMREAD_COLOR)


def gpu_ingest_enabled() -> bool:
    return os.getenv('MELF_INGEST', 'gpu').lower() != 'host'


def read_stream(filename: str) -> Optional[bytes]:
    """The file's bytes, or None where ``cv2.imread`` would return None for I/O reasons."""
    try:
        with open(filename, 'rb') as fp:
            return fp.read()
    except OSError:
        return None


def exif_orientation(data: bytes) -> int:
    """EXIF orientation tag (0x0112) of a JPEG stream, 1 when absent.  ``cv2.imread``
    rotates by it; the ingest kernels do not, so anything but 1 stays on the host."""
    pos = 2
    while pos + 4 <= len(data) and data[pos] == 0xFF:
        marker = data[pos + 1]
        if marker in (0xDA, 0xD9):
            break
        size = int.from_bytes(data[pos + 2:pos + 4], 'big')
        seg = data[pos + 4:pos + 2 + size]
        if marker == 0xE1 and seg[:6] == b'Exif\0\0' and len(seg) >= 14:
            tiff = seg[6:]
            order = 'little' if tiff[:2] == b'II' else 'big'
            ifd = int.from_bytes(tiff[4:8], order)
            if ifd + 2 > len(tiff):
                return 1
            for k in range(int.from_bytes(tiff[ifd:ifd + 2], order)):
                entry = tiff[ifd + 2 + 12 * k:ifd + 14 + 12 * k]
                if len(entry) == 12 and int.from_bytes(entry[:2], order) == 0x0112:
                    return int.from_bytes(entry[8:10], order)
            return 1
        pos += 2 + size
    return 1


def gpu_ingest_geometry(data: bytes) -> Optional[Tuple[int, ...]]:
    """(h, w, components, h_samp, v_samp) when ``melf_read_jpeg_batch`` can take the
    stream as it is, else None (the file then goes through ``cv2.imread``)."""
    from . import _engine
    if data[:2] != b'\xff\xd8':
        return None
    if b'\xff\xd9' not in data[-64:]:
        # no end-of-image marker: a file cut short.  libjpeg zero-fills what is missing in its own
        # way (jdhuff.c insufficient_data); let cv2.imread decide what the reference sees
        return None
    try:
        info = _engine.jpeg_info(data)
    except _engine.EngineError:
        return None
    if exif_orientation(data) != 1:
        return None
    return (info['h'], info['w'], info['components'], info['h_samp'], info['v_samp'])


class ImageFile:
    def __init__(self, filename: str, params: _Params,
                 bgr_image: Optional[Image] = None) -> None:
        self.filename = filename
        self.params = params
        self.bgr_image = bgr_image

    # -- host side ----------------------------------------------------------------------
    def load_frame(self) -> Tuple[Image, bool]:
        """(frame, needs_crop): an injected ``bgr_image`` is used as is, un-cropped
        (``_image.py:47-48``); a decoded file still has to be cut to ``meter_rect``."""
        if self.bgr_image is not None:
            return (self.bgr_image, False)
        img = _imread(self.filename)
        if img is None:
            raise ImageLoadingError(self.filename)
        return (img, True)

    def get_bgr_image(self) -> Image:
        (frame, needs_crop) = self.load_frame()
        if not needs_crop:
            return frame
        ((x0, y0), (x1, y1)) = self.params.meter_rect
        return frame[y0:y1, x0:x1]

    # -- GPU side -----------------------------------------------------------------------
    def get_hls_image(self) -> Image:
        from ._utils import convert_to_hls
        return convert_to_hls(self.get_bgr_image(), self.params.hue_shift)

    def find_dials(self) -> TemplateMatchResult:
        """``_find_dials`` (``_image.py:57-66``): TM_CCOEFF arg-max on the lightness plane."""
        from ._utils import match_template_bgr
        template = get_dials_template(self.params)
        match = match_template_bgr(self.get_bgr_image(), template, self.params)
        if match.max_val < self.params.dials_match_threshold:
            raise DialsNotFoundError(
                self.filename, extra_info={'match val': match.max_val})
        return match

    def _find_dials(self, img_hls: Image) -> TemplateMatchResult:
        """The reference's signature (``_image.py:57-66``): locate on the L plane of an HLS image."""
        from ._utils import match_template
        match = match_template(numpy.ascontiguousarray(img_hls[:, :, 1]), get_dials_template(self.params))
        if match.max_val < self.params.dials_match_threshold:
            raise DialsNotFoundError(
                self.filename, extra_info={'match val': match.max_val})
        return match

    def get_bgr_image_t(self) -> Image:
        """The BGR image translated so that the matched dials rect lands on (30, 116)
        (``_image.py:34-44``: ``cv2.warpAffine`` with an integer translation and a zero border is
        an exact pixel copy; the batched form is ``melf_average_translated``)."""
        bgr_image = self.get_bgr_image()
        (x0, y0) = self.find_dials().rect.top_left
        (dx, dy) = (30 - x0, 116 - y0)
        (h, w) = bgr_image.shape[:2]
        out = numpy.zeros_like(bgr_image)
        (sx0, sx1) = (max(0, -dx), min(w, w - dx))
        (sy0, sy1) = (max(0, -dy), min(h, h - dy))
        if sx1 > sx0 and sy1 > sy0:
            out[sy0 + dy:sy1 + dy, sx0 + dx:sx1 + dx] = bgr_image[sy0:sy1, sx0:sx1]
        return out

    def get_dials_hls(self) -> Image:
        match = self.find_dials()
        ((x0, y0), (x1, y1)) = match.rect
        return self.get_hls_image()[y0:y1, x0:x1]


_template_cache: Dict[int, Tuple[_Params, Image]] = {}


def get_dials_template(params: _Params) -> Image:
    entry = _template_cache.get(id(params))
    if entry is None or entry[0] is not params:
        template = _imread(params.dials_file, grayscale=True)
        if template is None:
            raise IOError("Cannot read dials template: {}".format(params.dials_file))
        entry = (params, template)
        _template_cache[id(params)] = entry
    assert entry[1].shape == params.dials_template_size
    return entry[1]


def as_rect(top_left: Tuple[int, int], size_wh: Tuple[int, int]) -> Rect:
    return Rect(top_left, (top_left[0] + size_wh[0], top_left[1] + size_wh[1]))


def stack_frames(frames) -> numpy.ndarray:
    return numpy.stack([numpy.ascontiguousarray(f) for f in frames])
