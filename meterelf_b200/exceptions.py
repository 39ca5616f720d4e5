"""Error taxonomy of the dial-reading path.

Mirrors the public behaviour of the reference's ``meterelf/exceptions.py:4-52`` (class
names, base classes, default messages and the ``get_message`` format), because the CLI
prints ``error.get_message()`` and callers catch these classes.  The per-frame status
codes returned by the C-ABI (``include/meterelf_b200.h``: ``MELF_ST_*``) are mapped onto
these classes in ``_reading.py``.
"""
from typing import Any, Dict, Optional


class ImageProcessingError(Exception):
    default_message = "Unable to process image"

    def __init__(self, filename: str = '', message: Optional[str] = None,
                 extra_info: Optional[Dict[str, Any]] = None) -> None:
        super().__init__()
        self.filename = filename
        self.message = message if message else self.default_message
        self.extra_info = extra_info

    def get_message(self, *, with_filename: bool = False,
                    with_extra_info: bool = True) -> str:
        parts = [self.message]
        if with_filename and self.filename:
            parts.append(' from file: {}'.format(self.filename))
        if with_extra_info and self.extra_info:
            details = ', '.join(
                '{} = {}'.format(key, val) for (key, val) in self.extra_info.items())
            if details:
                parts.append(' ({})'.format(details))
        return ''.join(parts)

    def __str__(self) -> str:
        return self.get_message(with_filename=True, with_extra_info=True)


class ImageLoadingError(ImageProcessingError, IOError):
    default_message = "Unable to load image"


class ImageAnalyzingError(ImageProcessingError, ValueError):
    default_message = "Failed to analyze image"


class DialsNotFoundError(ImageAnalyzingError):
    default_message = "Dials not found"


class DialAngleDeterminingError(ImageAnalyzingError):
    default_message = "Cannot determine angle of a dial"


class NeedleContoursNotFoundError(ImageAnalyzingError):
    default_message = "Cannot find needle contours of a dial"
