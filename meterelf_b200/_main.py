"""Command line front-end; output format identical to the reference (``_main.py:8-22``)."""
import sys
from typing import Sequence

from . import _debug
from ._api import get_meter_values


def format_line(data) -> str:
    """One stdout line per image: ``<file>: <value %07.3f>`` or ``<file>: UNKNOWN <msg>``.

    Like the reference, a falsy value (None or exactly 0.0) prints as an empty string.
    """
    text = data.filename + ': '
    if data.value:
        text += '{:07.3f}'.format(data.value)
    if data.error:
        text += 'UNKNOWN {}'.format(data.error.get_message())
    if _debug.DEBUG:
        text += ' {!r}'.format(data.meter_values)
    return text


def main(argv: Sequence[str] = sys.argv) -> None:
    if len(argv) < 2:
        raise SystemExit('Usage: {} PARAMETERS_FILE [IMAGE_FILE...]'.format(
            argv[0] if argv else 'meterelf'))
    for data in get_meter_values(argv[1], argv[2:]):
        print(format_line(data))  # noqa
