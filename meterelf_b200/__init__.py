"""meterelf_b200 -- B200-native implementation of meterelf's per-frame dial-reading path.

Public surface = the reference's (``meterelf/__init__.py:1-6``): ``get_meter_values`` and
``MeterImageData``.  The arithmetic runs in hand-written sm_100a CUDA kernels behind the
C ABI in ``include/meterelf_b200.h``; there is no CPU fallback.
"""
from ._api import MeterImageData, get_meter_values

__all__ = ['MeterImageData', 'get_meter_values']
