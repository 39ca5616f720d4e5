"""``DEBUG`` environment switch (reference: ``_debug.py:1-14``).

The CLI appends the meter_values dict and per-image errors are re-raised, as in the reference.
The reference's HighGUI windows (``cv2.imshow``) cannot exist on this box; with
``MELF_DEBUG_DIR=<dir>`` the same pictures are written as PNG files instead (``_overlay.py``).
"""
import os

_OFF_WORDS = {'0', 'no', 'off', 'false'}

DEBUG = set(
    word for word in os.getenv('DEBUG', '').replace(',', ' ').split()
    if word.lower() not in _OFF_WORDS)

if 'all' in DEBUG:
    DEBUG = {'masks'}


def reraise_if_debug_on() -> None:
    """Call inside an ``except`` block: re-raises the active exception in debug mode."""
    if DEBUG:
        raise
