"""Reading the dials (reference: ``_reading.py:19-182``), on the GPU, in batches.

``get_meter_value(imgf)`` keeps the reference's signature and error behaviour; the work
is done by ``read_image_files`` which maps the C ABI's per-frame status codes back onto
the reference's exception classes, messages and ``extra_info`` dicts.
"""
import re
import threading
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy

from . import _engine, _overlay
from ._dial_data import get_dial_data
from ._image import ImageFile, get_dials_template, gpu_ingest_enabled, stack_frames
from ._params import Params as _Params
from .exceptions import (
    DialAngleDeterminingError, DialsNotFoundError, ImageProcessingError,
    NeedleContoursNotFoundError)

_FULL_FRAME = (0, 0, 2 ** 30, 2 ** 30)
_engines: Dict[Tuple[int, bool, int], Tuple[_Params, _engine.Engine]] = {}
_engines_lock = threading.Lock()


def get_engine(params: _Params, cropped_input: bool = False, device: int = 0) -> _engine.Engine:
    """The ``melf_ctx`` for a params object (replaces the reference's per-params caches).

    ``cropped_input``: frames are already cut to ``meter_rect`` (injected ``bgr_image``)."""
    key = (id(params), cropped_input, device)
    with _engines_lock:
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
    return records_to_python(numpy.asarray(record).reshape(1), names, [filename])[0]


def records_to_python(records: numpy.ndarray, names: Sequence[str], filenames: Sequence[str],
                      ) -> List[Tuple[Dict[str, float], Optional[ImageProcessingError]]]:
    """``melf_result`` records -> [(meter_values dict, error)]; the fields leave numpy once per batch,
    not once per record (the per-file cost of the Python layer is what bounds get_meter_values)."""
    n_dials = len(names)
    status = records['status'].tolist()
    positions = records['dial_pos'][:, :n_dials].tolist()
    value = records['value'].tolist()
    out: List[Any] = []
    for (i, st) in enumerate(status):
        if st == _engine.ST_OK:
            if n_dials != 4:
                # determine_value_by_dial_positions asserts exactly four dials (_reading.py:166)
                raise AssertionError('exactly 4 dials are needed to compute a value')
            values = dict(zip(names, positions[i]))
            if value[i] == value[i]:
                values['value'] = value[i]
            out.append((values, None))
        elif st == _engine.ST_DIALS_NOT_FOUND:
            out.append(({}, DialsNotFoundError(
                filenames[i], extra_info={'match val': float(records['max_val'][i])})))
        elif st == _engine.ST_NO_CONTOURS:
            # raised without a filename in the reference (_reading.py:138)
            out.append(({}, NeedleContoursNotFoundError(
                extra_info={'dial': names[int(records['failed_dial'][i])]})))
        elif st == _engine.ST_ANGLE_UNDETERMINED:
            mask = int(records['unreadable_mask'][i])
            unreadable = [name for (k, name) in enumerate(names) if (mask >> k) & 1]
            out.append(({}, DialAngleDeterminingError(
                filenames[i], extra_info={'unreadable dials': ', '.join(unreadable)})))
        else:
            raise RuntimeError('unknown status {} from the CUDA library'.format(st))
    return out


class Deferred:
    """An exception the reference would raise when it REACHES a file (``cv2.error`` from
    ``cv2.matchTemplate`` on a frame smaller than the template): kept in that file's slot so that
    ``get_meter_values`` still yields the files before it, then raises."""

    def __init__(self, error: Exception) -> None:
        self.error = error


def read_image_files(params: _Params, imgfs: Sequence[ImageFile], device: int = 0,
                     ) -> List[Any]:
    """[(meter_values, error) or Deferred] per image.  Files are read and header-probed as ONE batch
    on the host side of the C ABI (``melf_read_files`` + ``melf_jpeg_probe_batch``); baseline JPEG
    streams the ingest kernels support go to the GPU as bytes, grouped by geometry
    (``melf_read_jpeg_batch``); anything else is decoded on the host as in the reference
    (``cv2.imread``), grouped by frame shape and read in one GPU batch per group."""
    out: List[Any] = [None] * len(imgfs)
    debug_dir = _overlay.overlay_dir()      # debug pictures need the decoded pixels on the host
    host_decode: List[int] = []
    from_file = [i for (i, imgf) in enumerate(imgfs) if imgf.bgr_image is None]
    if from_file and gpu_ingest_enabled() and debug_dir is None:
        (blob, offsets, readable) = _engine.read_files([imgfs[i].filename for i in from_file])
        try:
            _read_jpeg_groups(params, imgfs, from_file, blob, offsets, readable, host_decode, out, device)
        finally:
            _engine.release_blob(blob)
    else:
        host_decode.extend(from_file)
    _read_on_host(params, imgfs, host_decode, out, device, debug_dir)
    return out


def _read_jpeg_groups(params: _Params, imgfs: Sequence[ImageFile], from_file: List[int], blob: numpy.ndarray,
                      offsets: numpy.ndarray, readable: numpy.ndarray, host_decode: List[int], out: List[Any],
                      device: int) -> None:
    """The files whose bytes are in ``blob``: baseline JPEG streams go to the GPU grouped by geometry, the
    indices of everything else are appended to ``host_decode``."""
    info = _engine.jpeg_probe_batch(blob, offsets)
    # the GPU decoder takes a stream when the probe accepts it, cv2.imread would not rotate it (EXIF)
    # and it has its end-of-image marker (a file cut short: let cv2.imread decide what the reference sees)
    good = readable & (info[:, 7] == 0) & (info[:, 5] == 1) & (info[:, 6] == 1)
    host_decode.extend(from_file[k] for k in numpy.flatnonzero(~good).tolist())
    which = numpy.flatnonzero(good)
    if len(which):
        (geoms, inverse) = numpy.unique(info[which, :5], axis=0, return_inverse=True)
        inverse = inverse.reshape(-1)
        engine = get_engine(params, cropped_input=False, device=device)
        for g in range(len(geoms)):
            members = which[inverse == g]
            if len(members) == len(from_file):
                (sub_blob, sub_offsets) = (blob, offsets)
            else:
                parts = [blob[int(offsets[k]):int(offsets[k + 1])] for k in members.tolist()]
                sub_offsets = numpy.zeros(len(parts) + 1, numpy.uint64)
                numpy.cumsum([len(p) for p in parts], out=sub_offsets[1:])
                sub_blob = numpy.concatenate(parts + [numpy.zeros(16, numpy.uint8)])
            indices = [from_file[k] for k in members.tolist()]
            records = None
            for _attempt in range(9):
                try:
                    records = engine.read_jpeg_blob(sub_blob, sub_offsets)
                    break
                except _engine.EngineError as error:
                    if error.code == _engine.ERR_TOO_SMALL:
                        for index in indices:
                            out[index] = Deferred(ValueError(
                                'image smaller than the dials template: ' + imgfs[index].filename))
                        indices = []
                        break
                    if error.code not in (_engine.ERR_INVALID, _engine.ERR_UNSUPPORTED):
                        raise
                    # a stream the header probe accepted but the decoder rejects: cv2.imread decides what
                    # the reference would have seen for THAT file; the others stay on the GPU
                    bad = re.search(r'stream (\d+):', str(error))
                    if bad is None or not 0 <= int(bad.group(1)) < len(indices) or len(indices) == 1:
                        break
                    drop = int(bad.group(1))
                    host_decode.append(indices.pop(drop))
                    lengths = numpy.diff(sub_offsets).astype(numpy.int64)
                    keep = [k for k in range(len(lengths)) if k != drop]
                    parts = [sub_blob[int(sub_offsets[k]):int(sub_offsets[k + 1])] for k in keep]
                    sub_offsets = numpy.zeros(len(parts) + 1, numpy.uint64)
                    numpy.cumsum(lengths[keep], out=sub_offsets[1:])
                    sub_blob = numpy.concatenate(parts + [numpy.zeros(16, numpy.uint8)])
            if records is None:
                host_decode.extend(indices)
                continue
            converted = records_to_python(records, engine.names, [imgfs[i].filename for i in indices])
            for (index, item) in zip(indices, converted):
                out[index] = item


def _read_on_host(params: _Params, imgfs: Sequence[ImageFile], host_decode: List[int], out: List[Any], device: int,
                  debug_dir: Optional[str]) -> None:
    """Files decoded with ``cv2.imread`` as in the reference and injected images, one GPU batch per frame shape."""
    groups: Dict[Tuple[bool, Tuple[int, ...]], List[Tuple[int, numpy.ndarray]]] = {}
    host_decode.extend(i for (i, imgf) in enumerate(imgfs) if imgf.bgr_image is not None)
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
        indices = [index for (index, _frame) in members]
        try:
            records = engine.read_frames(stack_frames([frame for (_i, frame) in members]))
        except _engine.EngineError as error:
            if error.code != _engine.ERR_TOO_SMALL:
                raise
            # cv2.matchTemplate would raise cv2.error on these frames in the reference
            for index in indices:
                out[index] = Deferred(ValueError(
                    'image smaller than the dials template: ' + imgfs[index].filename))
            continue
        converted = records_to_python(records, engine.names, [imgfs[i].filename for i in indices])
        for (index, item) in zip(indices, converted):
            out[index] = item
        if debug_dir is not None:
            ((x0, y0), (x1, y1)) = params.meter_rect if needs_crop else ((0, 0), (None, None))
            crops = stack_frames([frame[y0:y1, x0:x1] for (_i, frame) in members])
            _overlay.write_overlays(
                params, get_engine(params, cropped_input=True, device=device), crops,
                [imgfs[index].filename for index in indices], debug_dir)


def get_meter_value(imgf: ImageFile) -> Dict[str, float]:
    (item,) = read_image_files(imgf.params, [imgf])
    if isinstance(item, Deferred):
        raise item.error
    (values, error) = item
    if error is not None:
        raise error
    return values
