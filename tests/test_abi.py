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


def test_the_stub_in_integration_md_has_the_library_s_struct_layout():
    """INTEGRATION.md shows the ctypes stub a maintainer of the reference would add; its structures must be the
    ones the library was compiled with (sizes and field offsets of melf_dial, melf_params, melf_result)."""
    import numpy as np
    path = os.path.join(os.path.dirname(HEADER), os.pardir, 'INTEGRATION.md')
    with open(path) as fp:
        text = fp.read()
    code = []
    for name in ('_Dial', '_Params'):
        m = re.search(r'^class %s\(C\.Structure\):.*?\n(?=\n)' % name, text, flags=re.S | re.M)
        assert m, name
        code.append(m.group(0))
    m = re.search(r'^RESULT = np\.dtype\(.*?align=True\)', text, flags=re.S | re.M)
    assert m
    code.append(m.group(0))
    scope = {'C': ctypes, 'np': np}
    exec('\n'.join(code), scope)
    for (stub, ours) in ((scope['_Dial'], _engine.CDial), (scope['_Params'], _engine.CParams)):
        assert ctypes.sizeof(stub) == ctypes.sizeof(ours)
        assert [(f[0], getattr(stub, f[0]).offset) for f in stub._fields_] == \
               [(f[0], getattr(ours, f[0]).offset) for f in ours._fields_]
    assert scope['RESULT'].itemsize == _engine.RESULT_DTYPE.itemsize
    assert {k: v[1] for (k, v) in scope['RESULT'].fields.items()} == \
           {k: v[1] for (k, v) in _engine.RESULT_DTYPE.fields.items()}


def test_header_compiles_as_plain_c_and_agrees_with_the_binding(tmp_path):
    """include/meterelf_b200.h is a C header (C99, -pedantic clean): what gcc lays out is what the Python
    binding assumes."""
    import shutil
    import subprocess
    gcc = shutil.which('gcc')
    if gcc is None:
        import pytest
        pytest.skip('no gcc')
    src = tmp_path / 'layout.c'
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "meterelf_b200.h"\n'
        'int main(void) { printf("%zu %zu %zu %zu %zu %d\\n", sizeof(melf_dial), sizeof(melf_params), '
        'sizeof(melf_result), offsetof(melf_params, dials), offsetof(melf_result, dial_pos), MELF_MAX_DIALS); '
        'return 0; }\n')
    exe = tmp_path / 'layout'
    subprocess.run([gcc, '-std=c99', '-Wall', '-Wextra', '-pedantic', '-Werror', '-I', os.path.dirname(HEADER),
                    str(src), '-o', str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE).stdout.split()]
    assert got == [ctypes.sizeof(_engine.CDial), ctypes.sizeof(_engine.CParams), _engine.RESULT_DTYPE.itemsize,
                   _engine.CParams.dials.offset, _engine.RESULT_DTYPE.fields['dial_pos'][1], _engine.MAX_DIALS]

