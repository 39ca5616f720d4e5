"""Reading the dials (reference: ``_reading.py:19-182``), on the GPU, in batches.

``get_meter_value(imgf)`` keeps the reference's signature and error behaviour; the work
is done by ``read_image_files`` which maps the C ABI's per-frame status codes back onto
the reference's exception classes, messages and ``extra_info`` dicts.
"""
import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy

from . import _engine
from ._dial_data import get_dial_data
from ._image import ImageFile, get_dials_template, stack_frames
from ._params import Params as _Params
from .exceptions import (
    DialAngleDeterminingError, DialsNotFoundError, ImageProcessingError,
    NeedleContoursNotFoundError)

_FULL_FRAME = (0, 0, 2 ** 30, 2 ** 30)
_engines: Dict[Tuple[int, bool, int], Tuple[_Params, _engine.Engine]] = {}


def get_engine(params: _Params, cropped_input: bool = False, device: int = 0) -> _engine.Engine:
    """The ``melf_ctx`` for a params object (replaces the reference's per-params caches).

    ``cropped_input``: frames are already cut to ``meter_rect`` (injected ``bgr_image``)."""
    key = (id(params), cropped_input, device)
    entry = _engines.get(key)
    if entry is None or entry[0] is not params:
        dial_data = get_dial_data(params)
        rect = _FULL_FRAME if cropped_input else (
            params.meter_rect.top_left + params.meter_rect.bottom_right)
        (cparams, keep) = _engine.build_cparams(
            get_dials_template(params), rect, params.hue_shift,
            params.dials_match_threshold, _engine.dials_from_params(params, dial_data))
        entry = (params, _engine.Engine(cparams, keep, list(dial_data.keys()), device))
        _engines[key] = entry
    return entry[1]


def result_to_python(record: Any, names: Sequence[str], filename: str,
                     ) -> Tuple[Dict[str, float], Optional[ImageProcessingError]]:
    """One ``melf_result`` -> (meter_values dict, error), as get_meter_value would give."""
    status = int(record['status'])
    if status == _engine.ST_OK:
        values = {name: float(record['dial_pos'][i]) for (i, name) in enumerate(names)}
        value = float(record['value'])
        if len(names) != 4:
            # determine_value_by_dial_positions asserts exactly four dials (_reading.py:166)
            raise AssertionError('exactly 4 dials are needed to compute a value')
        if not math.isnan(value):
            values['value'] = value
        return (values, None)
    if status == _engine.ST_DIALS_NOT_FOUND:
        return ({}, DialsNotFoundError(
            filename, extra_info={'match val': float(record['max_val'])}))
    if status == _engine.ST_NO_CONTOURS:
        # raised without a filename in the reference (_reading.py:138)
        return ({}, NeedleContoursNotFoundError(
            extra_info={'dial': names[int(record['failed_dial'])]}))
    if status == _engine.ST_ANGLE_UNDETERMINED:
        unreadable = [name for (i, name) in enumerate(names)
                      if (int(record['unreadable_mask']) >> i) & 1]
        return ({}, DialAngleDeterminingError(
            filename, extra_info={'unreadable dials': ', '.join(unreadable)}))
    raise RuntimeError('unknown status {} from the CUDA library'.format(status))


def read_image_files(params: _Params, imgfs: Sequence[ImageFile], device: int = 0,
                     ) -> List[Tuple[Dict[str, float], Optional[ImageProcessingError]]]:
    """Decode on the host, group by frame shape, read each group in one GPU batch."""
    out: List[Any] = [None] * len(imgfs)
    groups: Dict[Tuple[bool, Tuple[int, ...]], List[Tuple[int, numpy.ndarray]]] = {}
    for (index, imgf) in enumerate(imgfs):
        try:
            (frame, needs_crop) = imgf.load_frame()
        except ImageProcessingError as error:
            out[index] = ({}, error)
            continue
        groups.setdefault((needs_crop, frame.shape), []).append((index, frame))
    for ((needs_crop, _shape), members) in groups.items():
        engine = get_engine(params, cropped_input=not needs_crop, device=device)
        try:
            records = engine.read_frames(stack_frames([frame for (_i, frame) in members]))
        except _engine.EngineError as error:
            if error.code != _engine.ERR_TOO_SMALL:
                raise
            # cv2.matchTemplate would raise cv2.error on these frames in the reference
            raise ValueError(
                'image smaller than the dials template: ' + imgfs[members[0][0]].filename)
        for ((index, _frame), record) in zip(members, records):
            out[index] = result_to_python(record, engine.names, imgfs[index].filename)
    return out


def get_meter_value(imgf: ImageFile) -> Dict[str, float]:
    ((values, error),) = read_image_files(imgf.params, [imgf])
    if error is not None:
        raise error
    return values
