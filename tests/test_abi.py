"""The C-ABI library loads without a GPU and exports every function that
include/meterelf_b200.h declares (the drop-in boundary)."""
import ctypes
import os
import re

from meterelf_b200 import _engine

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                      'include', 'meterelf_b200.h')


def declared_functions():
    with open(HEADER) as fp:
        text = re.sub(r'/\*.*?\*/', '', fp.read(), flags=re.S)
    text = re.sub(r'^\s*#.*$', '', text, flags=re.M)
    # prototypes: "<type> [*]name(args);" outside of struct bodies
    text = re.sub(r'\{[^{}]*\}', '{}', text)
    return sorted(set(re.findall(r'\b([a-z_][a-z0-9_]*)\s*\([^;{}]*\)\s*;', text)))


def test_header_declares_the_expected_entry_points():
    names = declared_functions()
    for must in ('melf_create', 'melf_destroy', 'melf_read_batch', 'melf_read_batch_async',
                 'melf_convert_hls', 'melf_match_template', 'melf_last_error'):
        assert must in names
    assert all(n.startswith('melf_') for n in names), names


def test_library_exports_every_declared_symbol():
    lib = _engine.load_library()                  # dlopen only; needs no GPU
    missing = []
    for name in declared_functions():
        try:
            getattr(lib, name)
        except AttributeError:
            missing.append(name)
    assert not missing, missing
    assert isinstance(lib, ctypes.CDLL)
    assert lib.melf_abi_version() == _engine.ABI_VERSION


def test_python_binding_covers_the_header():
    assert set(declared_functions()) <= set(_engine._SYMBOLS)
