"""``params.yml`` loader -- the schema is part of the drop-in contract.

Same keys, same strict type checks (a YAML int is not accepted where a float is
required and vice versa), same ``LoadError`` behaviour and the same ``Params``
attribute names as the reference (``_params.py:17-155``), so a params file written for
the reference loads unchanged.  ``dials_template_size`` is given as [w, h] in the file
and exposed as (h, w) (``_params.py:136-138``).
"""
import os
from typing import Any, Dict, List, Mapping, Optional, Set

import yaml

from ._colors import HlsColor
from ._types import DialCenter, Rect, Size


class LoadError(Exception):
    pass


def _typed(data: Mapping[Any, Any], key: str, kind: type) -> Any:
    value = data[key]           # a missing key is a KeyError, as in the reference
    if not isinstance(value, kind):
        raise LoadError('{} is not {}'.format(key, kind.__name__))
    return value


def _typed_list(data: Mapping[Any, Any], key: str, kind: type,
                length: Optional[int] = None) -> List[Any]:
    items = _typed(data, key, list)
    for (index, item) in enumerate(items):
        if not isinstance(item, kind):
            raise LoadError('Item {} in {} is not {}'.format(index, key, kind.__name__))
    if length is not None and len(items) != length:
        raise LoadError('{} must have exactly {} items'.format(key, length))
    return items


def _hls(data: Mapping[Any, Any], key: str) -> HlsColor:
    sub = data[key]
    return HlsColor(_typed(sub, 'h', int), _typed(sub, 'l', int), _typed(sub, 's', int))


class _Needle:
    def __init__(self, data: Mapping[Any, Any]) -> None:
        self.name: str = _typed(data, 'name', str)
        self.color_range = _hls(data, 'color_range')
        self.dist_from_center: int = _typed(data, 'dist_from_center', int)
        self.circle_thickness: int = _typed(data, 'circle_thickness', int)
        self.angle_of_zero: float = _typed(data, 'angle_of_zero', float)
        (x, y) = _typed_list(data, 'center', float, 2)
        self.center = (x, y)
        self.diameter: int = _typed(data, 'diameter', int)
        self.negative_momentum: bool = _typed(data, 'negative_momentum', bool)


class Params:
    def __init__(self, base_dir: str, data: Dict[Any, Any]) -> None:
        def in_base_dir(key: str) -> str:
            name = _typed(data, key, str)
            return os.path.join(base_dir, name) if base_dir else name

        self.image_glob: str = in_base_dir('image_glob')

        rect = data['meter_rect']
        (x0, y0) = _typed_list(rect, 'top_left', int, 2)
        (x1, y1) = _typed_list(rect, 'bottom_right', int, 2)
        self.meter_rect = Rect(top_left=(x0, y0), bottom_right=(x1, y1))

        self.dials_file: str = in_base_dir('dials_template')
        if not os.path.exists(self.dials_file):
            raise LoadError('File not found: {}'.format(self.dials_file))
        self.dials_match_threshold: int = _typed(
            data, 'dials_template_match_threshold', int)
        (tmpl_w, tmpl_h) = _typed_list(data, 'dials_template_size', int, 2)
        self.dials_template_size: Size = (tmpl_h, tmpl_w)

        self.hue_shift: int = _typed(data, 'hue_shift', int)
        self.needle_color = _hls(data, 'needle_color')
        self.needle_color_range = _hls(data, 'needle_color_range')

        needle_dicts = _typed_list(data, 'needle_data', dict)
        if not needle_dicts:
            raise LoadError('Must have data of at least one needle')
        needles = [_Needle(item) for item in needle_dicts]

        # dicts keep YAML order: that is the dial evaluation order (_reading.py:28)
        self.dial_color_range: Dict[str, HlsColor] = {
            n.name: n.color_range for n in needles}
        self.needle_dists_from_dial_center: Dict[str, int] = {
            n.name: n.dist_from_center for n in needles}
        self.needle_circle_mask_thickness: Dict[str, int] = {
            n.name: n.circle_thickness for n in needles}
        self.needle_angles_of_zero: Dict[str, float] = {   # degrees
            n.name: n.angle_of_zero for n in needles}
        self.negative_momentum_dials: Set[str] = {
            n.name for n in needles if n.negative_momentum}
        self.dial_centers: Dict[str, DialCenter] = {
            n.name: DialCenter(n.center, n.diameter) for n in needles}

    @classmethod
    def load(cls, filename: str) -> 'Params':
        try:
            with open(filename, 'rt') as stream:
                data = yaml.load(stream, Loader=yaml.SafeLoader)
        except Exception as error:
            raise LoadError(
                'Cannot load YAML data from {}'.format(filename)) from error
        if not isinstance(data, dict):
            raise LoadError('Not a valid parameters file: {}'.format(filename))
        return cls(os.path.dirname(filename), data)


def load(filename: str) -> Params:
    return Params.load(filename)
