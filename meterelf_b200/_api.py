"""Public API (reference: ``_api.py:9-33``), batched.

``get_meter_values`` keeps the reference's contract -- a lazy generator that yields one
``MeterImageData`` per filename, in input order, with per-image ``ImageProcessingError``
captured in ``.error`` (re-raised in DEBUG mode) -- but decodes ``batch_size`` files at a
time and reads them in one GPU batch.
"""
import itertools
from typing import Dict, Iterable, Iterator, NamedTuple, Optional

from . import _debug, _params
from ._image import ImageFile
from ._reading import Deferred, read_image_files
from .exceptions import ImageProcessingError

# files per GPU batch: several waves of the persistent locate kernel (148 CTAs x 2 frames) and one
# chunk of the JPEG pipeline (8192 streams); ~300 MB of JPEG bytes on the host for the sample files
DEFAULT_BATCH_SIZE = 8192


class MeterImageData(NamedTuple):
    filename: str
    value: Optional[float]
    error: Optional[ImageProcessingError]
    meter_values: Dict[str, float]


def get_meter_values(params_file: str, filenames: Iterable[str],
                     batch_size: int = DEFAULT_BATCH_SIZE) -> Iterator[MeterImageData]:
    params = _params.load(params_file)
    names = iter(filenames)
    while True:
        chunk = list(itertools.islice(names, batch_size))
        if not chunk:
            return
        imgfs = [ImageFile(filename, params) for filename in chunk]
        for (imgf, item) in zip(imgfs, read_image_files(params, imgfs)):
            if isinstance(item, Deferred):
                raise item.error            # where the reference's loop would have raised (_api.py:22-33)
            (meter_values, error) = item
            if error is not None and _debug.DEBUG:
                raise error
            yield MeterImageData(imgf.filename, meter_values.get('value'), error, meter_values)
