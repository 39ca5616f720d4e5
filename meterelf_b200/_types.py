"""Plain data contracts shared by the host modules (reference: ``_types.py:1-31``)."""
from typing import NamedTuple, Tuple

import numpy

Image = numpy.ndarray
Point = Tuple[int, int]
FloatPoint = Tuple[float, float]
Size = Tuple[int, int]          # (height, width), like ``Params.dials_template_size``


class Rect(NamedTuple):
    top_left: Point
    bottom_right: Point


class DialCenter(NamedTuple):
    center: FloatPoint
    diameter: int


class DialData(NamedTuple):
    name: str
    center: FloatPoint           # the *float* centre from params.yml
    mask: Image                  # full disc, u8 {0,255}, shape dials_template_size
    circle_mask: Image           # annulus the needle tip crosses, same shape


class TemplateMatchResult(NamedTuple):
    rect: Rect
    max_val: float
