"""HLS colour triple used by params.yml and the thresholding step.

Behavioural mirror of the reference's ``HlsColor`` (``_colors.py:6-50``): a length-3
uint8 ndarray subclass, so ``image + HlsColor(shift, 0, 0)`` is numpy's wrapping u8 add,
and ``get_range`` clamps to [0, 255] per channel WITHOUT wrapping the hue (that is what
``hue_shift`` in params.yml is for).
"""
from typing import Tuple

import numpy


class HlsColor(numpy.ndarray):
    def __new__(cls, hue: int = 0, lightness: int = 0, saturation: int = 0) -> 'HlsColor':
        for channel in (hue, lightness, saturation):
            assert 0 <= channel < 256
        return numpy.array(
            [hue, lightness, saturation], dtype=numpy.uint8).view(cls)

    hue = property(lambda self: int(self[0]))
    lightness = property(lambda self: int(self[1]))
    saturation = property(lambda self: int(self[2]))

    def __repr__(self) -> str:
        return 'HlsColor({}, {}, {})'.format(self.hue, self.lightness, self.saturation)

    def get_range(self, color_range: 'HlsColor') -> Tuple['HlsColor', 'HlsColor']:
        mine = (self.hue, self.lightness, self.saturation)
        span = (color_range.hue, color_range.lightness, color_range.saturation)
        low = HlsColor(*(max(c - r, 0) for (c, r) in zip(mine, span)))
        high = HlsColor(*(min(c + r, 255) for (c, r) in zip(mine, span)))
        return (low, high)
