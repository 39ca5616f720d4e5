"""Run the UNMODIFIED reference (``/root/reference/meterelf``) in this container.

Test infrastructure only (never imported by the product).  The reference needs three
environment shims on a modern stack:

1. ``yaml.load`` without a Loader (``_params.py:22``) -> default to SafeLoader.
2. ``cv2.findContours`` returns 2 values on OpenCV 4, the reference unpacks 3
   (``_reading.py:132``) -> prepend a dummy.
3. ``cv2.ipp.setUseIPP(False)``: IPP's HLS conversion differs from the goldens.

``/root/reference`` exists only in the build container, so everything that uses this
module is either a generator script (``tests/golden/make_golden.py``) or a test that
skips itself when the reference is absent.
"""
import functools
import os
import sys

REFERENCE_ROOT = os.environ.get('MELF_REFERENCE_ROOT', '/root/reference')


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'meterelf'))


@functools.lru_cache(maxsize=1)
def load():
    """Import the reference package with the shims applied; returns the module."""
    if not available():
        raise RuntimeError('reference tree not present: ' + REFERENCE_ROOT)
    import cv2
    import yaml

    cv2.ipp.setUseIPP(False)
    cv2.setNumThreads(1)

    if not getattr(yaml.load, '_melf_shim', False):
        orig_load = yaml.load

        def load_with_default(stream, Loader=yaml.SafeLoader):
            return orig_load(stream, Loader=Loader)
        load_with_default._melf_shim = True
        yaml.load = load_with_default

    if not getattr(cv2.findContours, '_melf_shim', False):
        orig_fc = cv2.findContours

        def find_contours3(*args, **kwargs):
            res = orig_fc(*args, **kwargs)
            return (None,) + tuple(res) if len(res) == 2 else res
        find_contours3._melf_shim = True
        cv2.findContours = find_contours3

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import meterelf  # noqa: the reference
    assert os.path.realpath(meterelf.__file__).startswith(
        os.path.realpath(REFERENCE_ROOT)), meterelf.__file__
    return meterelf


def sample_dir(n: int) -> str:
    return os.path.join(REFERENCE_ROOT, 'sample-images%d' % n)


def sample_files(n: int):
    d = sample_dir(n)
    return sorted(f for f in os.listdir(d) if f.endswith('.jpg'))


def golden_stdout(n: int) -> str:
    with open(os.path.join(REFERENCE_ROOT, 'tests',
                           'sample-images%d_stdout.txt' % n)) as fp:
        return fp.read()
