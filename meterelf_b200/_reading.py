"""Reading the dials (reference: ``_reading.py:19-182``), on the GPU, in batches.

``get_meter_value(imgf)`` keeps the reference's signature and error behaviour; the work
is done by ``read_image_files`` which maps the C ABI's per-frame status codes back onto
the reference's exception classes, messages and ``extra_info`` dicts.
"""
import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy

from . import _engine, _overlay
from ._dial_data import get_dial_data
from ._image import (
    ImageFile, get_dials_template, gpu_ingest_enabled, gpu_ingest_geometry, read_stream,
    stack_frames)
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
    """Baseline JPEG files go to the GPU as bytes, grouped by geometry
    (``melf_read_jpeg_batch``); anything else is decoded on the host as in the reference,
    grouped by frame shape and read in one GPU batch per group."""
    out: List[Any] = [None] * len(imgfs)
    debug_dir = _overlay.overlay_dir()      # debug pictures need the decoded pixels on the host
    groups: Dict[Tuple[bool, Tuple[int, ...]], List[Tuple[int, numpy.ndarray]]] = {}
    streams: Dict[Tuple[int, ...], List[Tuple[int, bytes]]] = {}
    host_decode: List[int] = []
    for (index, imgf) in enumerate(imgfs):
        data = None
        if imgf.bgr_image is None and gpu_ingest_enabled() and debug_dir is None:
            data = read_stream(imgf.filename)
        geometry = gpu_ingest_geometry(data) if data else None
        if geometry is not None:
            streams.setdefault(geometry, []).append((index, data))
        else:
            host_decode.append(index)
    for members in streams.values():
        engine = get_engine(params, cropped_input=False, device=device)
        try:
            records = engine.read_jpeg_streams([data for (_i, data) in members])
        except _engine.EngineError as error:
            if error.code == _engine.ERR_TOO_SMALL:
                raise ValueError(
                    'image smaller than the dials template: ' + imgfs[members[0][0]].filename)
            if error.code not in (_engine.ERR_INVALID, _engine.ERR_UNSUPPORTED):
                raise
            # a stream the header probe accepted but the decoder rejects (e.g. cut short):
            # cv2.imread decides what the reference would have seen
            host_decode.extend(index for (index, _d) in members)
            continue
        for ((index, _d), record) in zip(members, records):
            out[index] = result_to_python(record, engine.names, imgfs[index].filename)
    for index in sorted(host_decode):
        imgf = imgfs[index]
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
        if debug_dir is not None:
            ((x0, y0), (x1, y1)) = params.meter_rect if needs_crop else ((0, 0), (None, None))
            crops = stack_frames([frame[y0:y1, x0:x1] for (_i, frame) in members])
            _overlay.write_overlays(
                params, get_engine(params, cropped_input=True, device=device), crops,
                [imgfs[index].filename for (index, _f) in members], debug_dir)
    return out


def get_meter_value(imgf: ImageFile) -> Dict[str, float]:
    ((values, error),) = read_image_files(imgf.params, [imgf])
    if error is not None:
        raise error
    return values
