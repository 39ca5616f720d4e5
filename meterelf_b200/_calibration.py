"""Offline calibration: dial centres from the average meter image (reference:
``_calibration.py:16-84``).

Same names and argument meaning as the reference.  The per-file work -- dial location
(``ImageFile._find_dials``), the translation of ``get_bgr_image_t``, the float64 running mean
and its denormalisation -- runs batched on the GPU (``melf_match_template`` +
``melf_average_translated``); the needle mask of the average image is the fused
convert + inRange operator.  ``cv2.findContours`` / ``cv2.fitEllipse`` over the four small
needle-hub contours stay on the host exactly as in the reference (offline, once per meter).
There is no CPU fallback for the GPU part.
"""
import glob
import random
from typing import Iterable, List, Sequence, Union

import numpy

from . import _engine
from ._image import ImageFile
from ._params import Params as _Params
from ._types import DialCenter, Image
from ._utils import _torch, get_mask_by_color
from .exceptions import DialsNotFoundError

#: where ``get_bgr_image_t`` moves the top-left corner of the matched dials rect (``_image.py:39-42``)
DIALS_ANCHOR = (30, 116)


def find_dial_centers(
        params: _Params,
        files: Union[int, Iterable[str]] = 255,
) -> List[DialCenter]:
    avg_meter = get_average_meter_image(params, get_files(params, files))
    return find_dial_centers_from_image(params, avg_meter)


def get_files(
        params: _Params,
        files: Union[int, Iterable[str]] = 255
) -> Iterable[str]:
    if isinstance(files, int):
        return random.sample(get_image_filenames(params), files)
    return files


def find_dial_centers_from_image(
        params: _Params,
        avg_meter: Image,
) -> List[DialCenter]:
    import cv2
    avg_meter_imgf = ImageFile('<average_image>', params, avg_meter)
    match = avg_meter_imgf.find_dials()
    ((x0, y0), (x1, y1)) = match.rect
    # mask of the whole image, then the dials rect: OpenCV's scalar-tail columns are
    # those of the full-width conversion, as in ImageFile.get_dials_hls
    needles_mask = numpy.ascontiguousarray(
        get_needles_mask_by_color(params, numpy.ascontiguousarray(avg_meter))[y0:y1, x0:x1])
    found = cv2.findContours(needles_mask, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_NONE)
    contours = found[-2]                      # OpenCV 3 returns (image, contours, hierarchy)
    dial_centers = []
    for contour in contours:
        (center, size, angle) = cv2.fitEllipse(contour)
        (height, width) = size
        diameter = (width + height) / 2.0
        if abs(height - width) / diameter > 0.2:
            raise ValueError('Needle center not circle enough')
        dial_centers.append(DialCenter(center, int(round(diameter))))
    return sorted(dial_centers, key=(lambda x: x.center[0]))


def get_average_meter_image(params: _Params, files: Iterable[str]) -> Image:
    names = list(files)
    if not names:
        raise ValueError("Cannot calculate average of empty sequence")
    crops = [ImageFile(fn, params).get_bgr_image() for fn in names]
    return average_of_crops(params, crops, names)


def average_of_crops(params: _Params, crops: Sequence[Image],
                     names: Union[Sequence[str], None] = None) -> Image:
    """``get_average_meter_image`` for already decoded + cropped frames of one size."""
    from ._reading import get_engine
    torch = _torch()
    frames = numpy.stack([numpy.ascontiguousarray(c) for c in crops])
    (n, h, w) = frames.shape[:3]
    engine = get_engine(params, cropped_input=True)
    stream = torch.cuda.current_stream().cuda_stream
    d_in = torch.from_numpy(frames).cuda()
    d_xy = torch.empty((n, 2), dtype=torch.int32, device='cuda')
    d_val = torch.empty((n,), dtype=torch.float32, device='cuda')
    engine.match_template_device(d_in.data_ptr(), 3, n, h, w, w * 3, h * w * 3, False,
                                 d_xy.data_ptr(), d_val.data_ptr(), stream=stream)
    vals = d_val.cpu().numpy()
    for (i, val) in enumerate(vals):          # get_bgr_image_t raises per file (_image.py:62-64)
        if float(val) < params.dials_match_threshold:
            raise DialsNotFoundError(names[i] if names else '<frame %d>' % i,
                                     extra_info={'match val': float(val)})
    d_out = torch.empty((h, w, 3), dtype=torch.uint8, device='cuda')
    _engine.average_translated_device(torch.cuda.current_device(), d_in.data_ptr(), n, h, w, w * 3,
                                      h * w * 3, d_xy.data_ptr(), DIALS_ANCHOR, d_out.data_ptr(), stream)
    return d_out.cpu().numpy()


def get_image_filenames(params: _Params) -> List[str]:
    return [
        path for path in glob.glob(params.image_glob)
        if all(bad_filename not in path for bad_filename in [
                '20180814021309-01-e01.jpg',
                '20180814021310-00-e02.jpg',
        ])
    ]


def get_needles_mask_by_color(params: _Params, bgr_image: Image) -> Image:
    """inRange around ``needle_color`` (``_calibration.py:82-84``); takes the BGR image because the
    conversion and the range test are one fused GPU pass."""
    return get_mask_by_color(bgr_image, params.hue_shift, params.needle_color,
                             params.needle_color_range)
