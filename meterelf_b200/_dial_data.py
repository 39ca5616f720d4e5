"""Per-params static dial masks (one-time host setup feeding the kernels).

Behavioural mirror of the reference's ``_dial_data.py:14-55``: for every dial a full
disc ``mask`` and an annulus ``circle_mask`` in dials-template coordinates.  The
reference draws them with ``cv2.circle`` + ``cv2.floodFill``; here both are restated in
numpy so the product has no OpenCV dependency on this step:

* ``cv2.circle(img, c, r, 255)`` (thickness 1, LINE_8) is OpenCV's integer midpoint
  circle: eight-way symmetric plotting with the error recurrence in ``_circle_points``.
* ``cv2.floodFill(img, mask, seed, 255)`` with default flags fills the 4-connected
  region of pixels equal to the seed's value.

``tests/test_host_dial_data.py`` pins both against cv2 (any radius, clipped circles)
and against the pixel counts SURVEY.md 8a-4 lists for the sample params.
"""
from typing import Dict, Iterator, Tuple

import numpy

from ._params import Params as _Params
from ._types import DialData, Image

_dial_data_cache: Dict[int, Tuple[_Params, Dict[str, DialData]]] = {}


def get_dial_data(params: _Params) -> Dict[str, DialData]:
    # keyed by id() like the reference (_dial_data.py:14-19), but the Params object is
    # kept alive in the cache entry so a recycled id can never alias another file
    entry = _dial_data_cache.get(id(params))
    if entry is None or entry[0] is not params:
        entry = (params, _build_dial_data(params))
        _dial_data_cache[id(params)] = entry
    return entry[1]


def py_round_int(value: float) -> int:
    """``int(round(x))`` -- Python's banker's rounding, as the reference uses."""
    return int(round(value))


def _circle_points(radius: int) -> Iterator[Tuple[int, int]]:
    """Offsets (dx, dy), dx >= dy >= 0, of one octant of OpenCV's 1-px circle."""
    err = 0
    dx = radius
    dy = 0
    plus = 1
    minus = 2 * radius - 1
    while dx >= dy:
        yield (dx, dy)
        dy += 1
        err += plus
        plus += 2
        if err > 0:
            err -= minus
            dx -= 1
            minus -= 2


def draw_circle(img: Image, center: Tuple[int, int], radius: int, value: int = 255) -> None:
    (height, width) = img.shape[:2]
    (cx, cy) = center
    for (dx, dy) in _circle_points(radius):
        for (px, py) in ((cx - dx, cy - dy), (cx + dx, cy - dy),
                         (cx - dx, cy + dy), (cx + dx, cy + dy),
                         (cx - dy, cy - dx), (cx + dy, cy - dx),
                         (cx - dy, cy + dx), (cx + dy, cy + dx)):
            if 0 <= px < width and 0 <= py < height:
                img[py, px] = value


def flood_fill(img: Image, seed: Tuple[int, int], value: int = 255) -> None:
    """4-connected fill of the region that has the seed pixel's value."""
    (height, width) = img.shape[:2]
    (sx, sy) = seed
    if not (0 <= sx < width and 0 <= sy < height):
        raise ValueError('flood fill seed outside of the image')   # cv2 raises too
    old = int(img[sy, sx])
    if old == value:
        return
    stack = [(sx, sy)]
    img[sy, sx] = value
    while stack:
        (x, y) = stack.pop()
        for (nx, ny) in ((x - 1, y), (x + 1, y), (x, y - 1), (x, y + 1)):
            if 0 <= nx < width and 0 <= ny < height and img[ny, nx] == old:
                img[ny, nx] = value
                stack.append((nx, ny))


def _build_dial_data(params: _Params) -> Dict[str, DialData]:
    result: Dict[str, DialData] = {}
    for (name, dial_center) in params.dial_centers.items():
        canvas = numpy.zeros(params.dials_template_size, dtype=numpy.uint8)
        dial_radius = py_round_int(dial_center.diameter / 2.0)
        center = (py_round_int(dial_center.center[0]),
                  py_round_int(dial_center.center[1]))

        inner_radius = dial_radius + params.needle_dists_from_dial_center[name]
        thickness = params.needle_circle_mask_thickness[name]
        for extra in (0, thickness - 1):
            draw_circle(canvas, center, inner_radius + extra)

        # annulus between the two outlines -> circle_mask
        flood_fill(canvas, (center[0] + inner_radius + 1, center[1]))
        circle_mask = canvas.copy()

        # plus everything inside the inner outline -> full disc
        flood_fill(canvas, center)
        result[name] = DialData(name, dial_center.center, canvas, circle_mask)
    return result
