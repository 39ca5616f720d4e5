"""Public API (reference: ``_api.py:9-33``), batched.

``get_meter_values`` keeps the reference's contract -- a lazy generator that yields one
``MeterImageData`` per filename, in input order, with per-image ``ImageProcessingError``
captured in ``.error`` (re-raised in DEBUG mode) -- but decodes ``batch_size`` files at a
time and reads them in one GPU batch.
"""
import collections
import itertools
from concurrent.futures import Future, ThreadPoolExecutor
from typing import Any, Dict, Iterable, Iterator, List, NamedTuple, Optional, Tuple

from . import _debug, _overlay, _params
from ._image import ImageFile
from ._reading import Deferred, read_image_files
from .exceptions import ImageProcessingError

# files per GPU batch: several waves of the persistent locate kernel (148 CTAs x 2 frames) and one
# chunk of the JPEG pipeline (8192 streams); ~300 MB of JPEG bytes on the host for the sample files
DEFAULT_BATCH_SIZE = 8192
# batches in flight on host threads (file reads + header probes of one under the GPU work of the other)
PREFETCH_BATCHES = 2


class MeterImageData(NamedTuple):
    filename: str
    value: Optional[float]
    error: Optional[ImageProcessingError]
    meter_values: Dict[str, float]


def get_meter_values(params_file: str, filenames: Iterable[str],
                     batch_size: int = DEFAULT_BATCH_SIZE) -> Iterator[MeterImageData]:
    params = _params.load(params_file)
    names = iter(filenames)

    def batches() -> Iterator[List[ImageFile]]:
        while True:
            chunk = list(itertools.islice(names, batch_size))
            if not chunk:
                return
            yield [ImageFile(filename, params) for filename in chunk]

    def results(imgfs: List[ImageFile], items: List[Any]) -> Iterator[MeterImageData]:
        for (imgf, item) in zip(imgfs, items):
            if isinstance(item, Deferred):
                raise item.error            # where the reference's loop would have raised (_api.py:22-33)
            (meter_values, error) = item
            if error is not None and _debug.DEBUG:
                raise error
            yield MeterImageData(imgf.filename, meter_values.get('value'), error, meter_values)

    if _debug.DEBUG or _overlay.overlay_dir() is not None:
        # debug mode stops at the first error and writes pictures: strictly one batch at a time
        for imgfs in batches():
            yield from results(imgfs, read_image_files(params, imgfs))
        return
    # Two batches ahead on two host threads: while the GPU decodes and reads batch k (one call at a time per
    # context, Engine._call_lock), the files of batch k + 1 are read and header-probed.  A batch is only
    # started once the consumer has asked for the first result of the one two before it.
    pool = ThreadPoolExecutor(max_workers=PREFETCH_BATCHES, thread_name_prefix='meterelf-batch')
    pending: 'collections.deque[Tuple[List[ImageFile], Future]]' = collections.deque()
    source = batches()
    try:
        while True:
            while len(pending) < PREFETCH_BATCHES:
                imgfs = next(source, None)
                if imgfs is None:
                    break
                pending.append((imgfs, pool.submit(read_image_files, params, imgfs)))
            if not pending:
                return
            (imgfs, future) = pending.popleft()
            yield from results(imgfs, future.result())
    finally:
        for (_imgfs, future) in pending:
            future.cancel()
        pool.shutdown(wait=True)
