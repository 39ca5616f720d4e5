"""Host-side static tables of the CUDA library, checked on the CPU (no GPU needed).

* the tensor-core TM_CCOEFF formulation (Toeplitz operand image + strip layout + M-tile
  plan, modelled in numpy by ``kernel_model.match_tc_model``) reproduces the direct
  correlation sum for the sample geometry and for ragged sizes;
* the per-dial tables (``melf_host_dial_tables``) agree with the masks they come from.
"""
import ctypes as C

import numpy as np
import pytest

import kernel_model
from helpers import golden_set
from meterelf_b200 import _engine


def _lib():
    return _engine.load_library()


def _toeplitz(tmpl):
    lib = _lib()
    lib.melf_host_match_tc_toeplitz.restype = C.c_longlong
    lib.melf_host_match_tc_toeplitz.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_longlong]
    (th, tw) = tmpl.shape
    need = lib.melf_host_match_tc_toeplitz(tmpl.ctypes.data, tw, th, None, 0)
    assert need == ((th + 1) // 2) * 32768
    buf = np.zeros(need, np.uint8)
    assert lib.melf_host_match_tc_toeplitz(tmpl.ctypes.data, tw, th, buf.ctypes.data, need) == need
    return buf.tobytes()


def _plan(res_w, res_h, tw, th):
    lib = _lib()
    plan = (C.c_int32 * 4)()
    assert lib.melf_host_match_tc_plan(res_w, res_h, tw, th, plan) == 0
    return tuple(plan)


def _direct(light, tmpl):
    (th, tw) = tmpl.shape
    (h, w) = light.shape
    win = np.lib.stride_tricks.sliding_window_view(light.astype(np.int64), (th, tw))
    return np.einsum('yxij,ij->yx', win, tmpl.astype(np.int64))


def test_match_tc_model_sample_geometry():
    gs = golden_set(1)
    crop = gs.subset()['crops'][3]
    light = (crop.max(axis=2).astype(np.int32) + crop.min(axis=2) + 1) // 2     # any u8 plane will do
    light = light.astype(np.uint8)
    tmpl = np.ascontiguousarray(gs.template)
    (th, tw) = tmpl.shape
    plan = _plan(light.shape[1] - tw + 1, light.shape[0] - th + 1, tw, th)
    assert plan == (1, 1, 132, 60)
    got = kernel_model.match_tc_model(light, _toeplitz(tmpl), tw, th, plan)
    assert np.array_equal(got, _direct(light, tmpl))


@pytest.mark.parametrize('crop_hw,tmpl_hw', [((135, 220), (119, 188)), ((40, 100), (7, 30)),
                                              ((300, 200), (20, 60)), ((150, 321), (12, 193)),
                                              ((128 + 8, 70), (9, 5))])
def test_match_tc_model_ragged_sizes(crop_hw, tmpl_hw):
    rng = np.random.default_rng(crop_hw[0] * 1000 + tmpl_hw[1])
    light = rng.integers(0, 256, size=crop_hw, dtype=np.uint8)
    tmpl = rng.integers(0, 256, size=tmpl_hw, dtype=np.uint8)
    (th, tw) = tmpl_hw
    plan = _plan(crop_hw[1] - tw + 1, crop_hw[0] - th + 1, tw, th)
    got = kernel_model.match_tc_model(light, _toeplitz(tmpl), tw, th, plan)
    assert np.array_equal(got, _direct(light, tmpl))


def test_match_tc_unsupported_templates_report_it():
    lib = _lib()
    lib.melf_host_match_tc_toeplitz.restype = C.c_longlong
    lib.melf_host_match_tc_toeplitz.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_longlong]
    assert lib.melf_host_match_tc_toeplitz(None, 194, 100, None, 0) == -1     # K = tw + 63 > 256
    assert lib.melf_host_match_tc_toeplitz(None, 100, 129, None, 0) == -1


@pytest.mark.parametrize('set_no', [1, 2])
def test_dial_tables_match_masks(set_no):
    gs = golden_set(set_no)
    lib = _lib()
    (cp, keep) = gs.cparams()
    for d in range(len(gs.names)):
        geom = (C.c_int32 * 6)()
        words = (C.c_uint64 * (2 + 3 * 64))()
        terms = (C.c_double * 128)()
        n_ring = C.c_int32()
        cap = 4096
        pos = (C.c_uint16 * cap)()
        ang = (C.c_double * cap)()
        sqd = (C.c_double * cap)()
        rc = lib.melf_host_dial_tables(C.byref(cp), d, geom, words, terms, C.byref(n_ring), pos, ang, sqd, cap)
        assert rc == 0
        (wx, wy, ww, wh, px, py) = list(geom)
        mask = gs.dials[d]['mask']
        ring = gs.dials[d]['circle_mask']
        (cx, cy) = gs.dials[d]['center']
        assert (px, py) == (int(cx) - 2, int(cy) - 2)
        disc = np.array(list(words)[2:66], np.uint64)
        ringw = np.array(list(words)[130:194], np.uint64)
        n_disc = sum(bin(int(w)).count('1') for w in disc)
        assert n_disc == int(np.count_nonzero(mask))
        assert sum(bin(int(w)).count('1') for w in ringw) == int(np.count_nonzero(ring))
        # every ring pixel with a defined angle is in the sorted table, ascending by (angle, sqd)
        n = n_ring.value
        assert 0 < n <= int(np.count_nonzero(ring))
        keys = [(ang[i], sqd[i]) for i in range(n)]
        assert keys == sorted(keys)
        for i in range(0, n, 7):
            (r, c) = (pos[i] >> 8, pos[i] & 255)
            (x, y) = (wx + c, wy + r)
            assert ring[y, x]
            a = kernel_model_angle(x - cx, y - cy)
            assert a == ang[i]
            assert sqd[i] == (x - cx) ** 2 + (y - cy) ** 2


def kernel_model_angle(x, y):
    import math
    if y == 0:
        return 0.25 if x > 0 else 0.75
    return (-(math.atan(x / y) / (2 * math.pi)) + (0.5 if y > 0 else 0.0)) % 1.0
