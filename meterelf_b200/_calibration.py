"""Offline calibration: where are the dial centres?  (reference: ``_calibration.py:16-84``)

The public names and what they take and return are the reference's.  How the work is done is not:

* per file, dial location + the translation of ``get_bgr_image_t`` + the float64 running mean + its
  denormalisation are ONE batched GPU job (``melf_match_template`` + ``melf_average_translated``);
* the needle-hub mask of the average image is the fused convert + inRange operator;
* the hub outlines are followed and fitted here, without OpenCV: ``outer_borders`` is border
  following after Suzuki & Abe (what ``cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_NONE)`` returns),
  ``fit_ellipse`` is the least-squares conic fit ``cv2.fitEllipse`` documents (general conic through
  the centred points, centre from the gradient equations, re-fit of the quadratic part, axes from
  its eigen-structure).  Both are checked against cv2 in ``tests/test_calibration.py``.

There is no CPU fallback for the GPU part.
"""
import glob
import math
import random
from typing import Iterable, List, Sequence, Tuple, Union

import numpy

from . import _engine
from ._image import ImageFile
from ._params import Params as _Params
from ._types import DialCenter, Image
from ._utils import _torch, get_mask_by_color_bgr
from .exceptions import DialsNotFoundError

#: where ``get_bgr_image_t`` moves the top-left corner of the matched dials rect (``_image.py:39-42``)
DIALS_ANCHOR = (30, 116)

#: sample files without a meter in them, never used for calibration (``_calibration.py:71-79``)
UNUSABLE_SAMPLES = ('20180814021309-01-e01.jpg', '20180814021310-00-e02.jpg')

#: largest relative difference of the two ellipse axes a needle hub may have (``_calibration.py:54-55``)
MAX_ECCENTRICITY = 0.2

# the 8 neighbours in clockwise order on the screen (y down), starting at west
_RING = ((-1, 0), (-1, -1), (0, -1), (1, -1), (1, 0), (1, 1), (0, 1), (-1, 1))


def find_dial_centers(params: _Params, files: Union[int, Iterable[str]] = 255) -> List[DialCenter]:
    """Centre and diameter of every needle hub, left to right."""
    return find_dial_centers_from_image(params, get_average_meter_image(params, get_files(params, files)))


def get_files(params: _Params, files: Union[int, Iterable[str]] = 255) -> Iterable[str]:
    """An explicit list is used as it is; a number draws that many of the params' image files."""
    if not isinstance(files, int):
        return files
    return random.sample(get_image_filenames(params), files)


def get_image_filenames(params: _Params) -> List[str]:
    usable = []
    for path in glob.glob(params.image_glob):
        if not any(name in path for name in UNUSABLE_SAMPLES):
            usable.append(path)
    return usable


def outer_borders(mask: Image) -> List[numpy.ndarray]:
    """Outer border of every 8-connected component of ``mask != 0`` as an (n, 2) array of (x, y),
    pixel by pixel in following order (a pixel the border passes twice is listed twice), found in
    raster order and returned newest first -- ``cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_NONE)``."""
    fg = numpy.pad(numpy.asarray(mask) != 0, 1)               # zero frame: no bounds checks below
    label = numpy.zeros(fg.shape, numpy.int32)
    # background reachable from the frame (4-connected): a component whose raster-first pixel has
    # enclosed background to its west sits in a hole of another component and is not external
    outside = numpy.zeros(fg.shape, bool)
    outside[0, 0] = True
    stack = [(0, 0)]
    (rows, cols) = fg.shape
    while stack:
        (cx, cy) = stack.pop()
        for (nx, ny) in ((cx - 1, cy), (cx + 1, cy), (cx, cy - 1), (cx, cy + 1)):
            if 0 <= nx < cols and 0 <= ny < rows and not fg[ny, nx] and not outside[ny, nx]:
                outside[ny, nx] = True
                stack.append((nx, ny))
    borders: List[numpy.ndarray] = []
    (ys, xs) = numpy.nonzero(fg)
    for (y0, x0) in zip(ys.tolist(), xs.tolist()):
        if label[y0, x0]:
            continue
        external = bool(outside[y0, x0 - 1])
        # mark the whole component so that its other pixels do not start a border of their own
        stack = [(x0, y0)]
        label[y0, x0] = 1
        while stack:
            (cx, cy) = stack.pop()
            for (dx, dy) in _RING:
                (nx, ny) = (cx + dx, cy + dy)
                if fg[ny, nx] and not label[ny, nx]:
                    label[ny, nx] = 1
                    stack.append((nx, ny))
        if external:
            borders.append(_follow_border(fg, x0, y0))
    return borders[::-1]


def _follow_border(fg: numpy.ndarray, x0: int, y0: int) -> numpy.ndarray:
    """Suzuki-Abe border following from the raster-first pixel of a component (its west
    neighbour is background).  ``fg`` carries a one-pixel zero frame; points come back unpadded."""
    points = [(x0 - 1, y0 - 1)]
    first = None
    for k in range(8):                                        # clockwise from west
        if fg[y0 + _RING[k][1], x0 + _RING[k][0]]:
            first = (x0 + _RING[k][0], y0 + _RING[k][1])
            break
    if first is None:
        return numpy.array(points, numpy.int32)               # a single pixel
    (prev, cur) = (first, (x0, y0))
    while True:
        start = _RING.index((prev[0] - cur[0], prev[1] - cur[1]))
        for t in range(1, 9):                                 # counter-clockwise, after `prev`
            (dx, dy) = _RING[(start - t) % 8]
            if fg[cur[1] + dy, cur[0] + dx]:
                nxt = (cur[0] + dx, cur[1] + dy)
                break
        if nxt == (x0, y0) and cur == first:
            break
        (prev, cur) = (cur, nxt)
        points.append((cur[0] - 1, cur[1] - 1))
    return numpy.array(points, numpy.int32)


def fit_ellipse(points: numpy.ndarray) -> Tuple[Tuple[float, float], Tuple[float, float], float]:
    """((cx, cy), (width, height), angle in degrees) of the least-squares ellipse through at least
    five points, with ``cv2.fitEllipse``'s conventions: float32 results, width <= height."""
    pts = numpy.asarray(points, numpy.float64).reshape(-1, 2)
    if len(pts) < 5:
        raise ValueError('at least five points are needed to fit an ellipse')
    mean = pts.astype(numpy.float32).mean(axis=0, dtype=numpy.float64).astype(numpy.float32).astype(numpy.float64)
    (x, y) = (pts[:, 0] - mean[0], pts[:, 1] - mean[1])
    # general conic  -A x^2 - B y^2 - C xy + D x + E y = 10000  through the centred points
    conic = numpy.linalg.lstsq(numpy.stack([-x * x, -y * y, -x * y, x, y], axis=1),
                               numpy.full(len(pts), 10000.0), rcond=None)[0]
    # its centre: where both partial derivatives vanish
    centre = numpy.linalg.lstsq(numpy.array([[2 * conic[0], conic[2]], [conic[2], 2 * conic[1]]]),
                                conic[3:5], rcond=None)[0]
    # quadratic part again, around that centre:  A u^2 + B v^2 + C uv = 1
    (u, v) = (x - centre[0], y - centre[1])
    (a, b, c) = numpy.linalg.lstsq(numpy.stack([u * u, v * v, u * v], axis=1), numpy.ones(len(pts)), rcond=None)[0]
    theta = -0.5 * math.atan2(c, b - a)
    t = c / math.sin(-2.0 * theta) if abs(c) > 1e-8 else b - a
    axes = []
    for q in (abs(a + b - t), abs(a + b + t)):
        axes.append(math.sqrt(2.0 / q) if q > 1e-8 else q)
    f32 = numpy.float32
    (width, height) = (float(f32(2 * axes[0])), float(f32(2 * axes[1])))
    angle = theta * 180.0 / math.pi
    if width > height:
        (width, height) = (height, width)
        angle += 90.0
    angle = float(f32(angle))
    if angle < -180:
        angle += 360
    if angle > 360:
        angle -= 360
    return ((float(f32(f32(centre[0]) + f32(mean[0]))), float(f32(f32(centre[1]) + f32(mean[1])))),
            (width, height), angle)


def find_dial_centers_from_image(params: _Params, avg_meter: Image) -> List[DialCenter]:
    located = ImageFile('<average_image>', params, avg_meter).find_dials()
    ((left, top), (right, bottom)) = located.rect
    # the mask is taken over the whole image and cut afterwards: OpenCV's scalar-tail columns
    #  are then those of the full-width conversion, as in ImageFile.get_dials_hls
    whole = get_needles_mask_by_color(params, numpy.ascontiguousarray(avg_meter))
    hubs = []
    for border in outer_borders(whole[top:bottom, left:right]):
        (centre, (width, height), _angle) = fit_ellipse(border)
        diameter = (width + height) / 2.0
        if abs(height - width) / diameter > MAX_ECCENTRICITY:
            raise ValueError('Needle center not circle enough')
        hubs.append(DialCenter(centre, int(round(diameter))))
    hubs.sort(key=lambda hub: hub.center[0])
    return hubs


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
    device = torch.cuda.current_device()
    frames = numpy.stack([numpy.ascontiguousarray(c) for c in crops])
    (n, h, w) = frames.shape[:3]
    engine = get_engine(params, cropped_input=True, device=device)
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
    _engine.average_translated_device(device, d_in.data_ptr(), n, h, w, w * 3,
                                      h * w * 3, d_xy.data_ptr(), DIALS_ANCHOR, d_out.data_ptr(), stream)
    return d_out.cpu().numpy()


def get_needles_mask_by_color(params: _Params, bgr_image: Image) -> Image:
    """inRange around ``needle_color`` (``_calibration.py:82-84``); takes the BGR image because the
    conversion and the range test are one fused GPU pass."""
    return get_mask_by_color_bgr(bgr_image, params.hue_shift, params.needle_color,
                                 params.needle_color_range)
