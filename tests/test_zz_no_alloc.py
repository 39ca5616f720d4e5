"""``melf_match_template_noalloc`` (include/meterelf_b200.h): the locate kernel on the library's static scratch.

No counterpart in the reference; the entry point was added for callers
that cannot call cudaMalloc.  Through the C ABI, against the existing ``melf_match_template`` on the same device
buffers: match location and ``max_val`` bit for bit (both are the same exact integer arithmetic), on calls in a
row that share the static array -- also with a different geometry in between, so that what one call leaves in
the array is what the next one finds.  Collected last on purpose (file name), passes alone and after the others.
"""
import numpy as np
import pytest

from helpers import golden_set
from meterelf_b200 import _engine

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def _both(eng, d_in, channels, n, h, w, use_rect=False):
    """(xy, val) of the existing entry point and of the no-alloc one, on the same device input."""
    torch = _torch()
    stream = torch.cuda.current_stream().cuda_stream
    out = []
    for call in (eng.match_template_device, eng.match_template_noalloc_device):
        d_xy = torch.full((n, 2), -7, dtype=torch.int32, device='cuda')
        d_val = torch.full((n,), -7.0, dtype=torch.float32, device='cuda')
        if call is eng.match_template_device:
            call(d_in.data_ptr(), channels, n, h, w, w * channels, h * w * channels, use_rect,
                 d_xy.data_ptr(), d_val.data_ptr(), 0, stream)
        else:
            call(d_in.data_ptr(), channels, n, h, w, w * channels, h * w * channels, use_rect,
                 d_xy.data_ptr(), d_val.data_ptr(), stream)
        torch.cuda.synchronize()
        out.append((d_xy.cpu().numpy(), d_val.cpu().numpy()))
    return out


def _assert_same(got, want, what):
    assert np.array_equal(got[0], want[0]), what
    assert np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32)), what      # float32, bit for bit


def test_noalloc_equals_match_template_on_calls_in_a_row():
    torch = _torch()
    gs = golden_set(1)
    eng = gs.engine(cropped_input=True)
    crops = gs.subset()['crops']
    (n, h, w) = crops.shape[:3]
    rng = np.random.default_rng(77)
    noise = rng.integers(0, 256, size=(300, h, w, 3), dtype=np.uint8)       # more than one wave of 2-frame passes
    noise[::3, 20:20 + gs.tmpl_h, 11:11 + gs.tmpl_w] = gs.template[None, :, :, None]
    gray = rng.integers(0, 256, size=(5, 200, 330), dtype=np.uint8)         # another geometry, 1 channel
    d_crops = torch.from_numpy(np.ascontiguousarray(crops)).cuda()
    d_noise = torch.from_numpy(noise).cuda()
    d_gray = torch.from_numpy(gray).cuda()
    # call 1 and call 2 in a row on the shared array, then a different geometry, then the first input again
    for (d_in, ch, nn, hh, ww, what) in ((d_crops, 3, n, h, w, 'sample crops'), (d_noise, 3, 300, h, w, 'noise'),
                                         (d_gray, 1, 5, 200, 330, 'gray, other geometry'),
                                         (d_crops, 3, n, h, w, 'sample crops again')):
        (want, got) = _both(eng, d_in, ch, nn, hh, ww)
        assert (want[0] != -7).all()
        _assert_same(got, want, what)
    # ... and it allocates nothing: device memory in use is the same before and after two more calls
    d_xy = torch.zeros((300, 2), dtype=torch.int32, device='cuda')
    d_val = torch.zeros((300,), dtype=torch.float32, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(2):
        eng.match_template_noalloc_device(d_noise.data_ptr(), 3, 300, h, w, w * 3, h * w * 3, False,
                                          d_xy.data_ptr(), d_val.data_ptr(), stream)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before


def test_noalloc_with_meter_rect_on_full_frames():
    torch = _torch()
    gs = golden_set(1)
    eng = gs.engine(cropped_input=False)
    rng = np.random.default_rng(5)
    frames = rng.integers(0, 256, size=(9, 480, 640, 3), dtype=np.uint8)
    d_in = torch.from_numpy(frames).cuda()
    for _ in range(2):
        (want, got) = _both(eng, d_in, 3, 9, 480, 640, use_rect=True)
        _assert_same(got, want, 'full frames')


def test_noalloc_rejects_what_does_not_fit_the_static_scratch():
    torch = _torch()
    gs = golden_set(1)
    eng = gs.engine(cropped_input=True)
    d_small = torch.zeros((1 << 20,), dtype=torch.uint8, device='cuda')
    d_xy = torch.zeros((4, 2), dtype=torch.int32, device='cuda')
    d_val = torch.zeros((4,), dtype=torch.float32, device='cuda')
    # the size check comes before anything touches the buffers
    with pytest.raises(_engine.EngineError) as err:
        eng.match_template_noalloc_device(d_small.data_ptr(), 3, _engine.NOALLOC_MAX_FRAMES + 1, 200, 300, 900, 180000,
                                          False, d_xy.data_ptr(), d_val.data_ptr(), 0)
    assert err.value.code == _engine.ERR_UNSUPPORTED and 'MELF_NOALLOC_MAX_FRAMES' in str(err.value)
    # a template the tensor-core kernel does not take (wider than 193 px): clear error, no fallback
    from meterelf_b200 import _utils
    wide = np.full((20, 200), 9, np.uint8)
    wide_eng = _utils._template_engine(wide, torch.cuda.current_device())
    with pytest.raises(_engine.EngineError) as err:
        wide_eng.match_template_noalloc_device(d_small.data_ptr(), 1, 1, 100, 400, 400, 40000, False,
                                               d_xy.data_ptr(), d_val.data_ptr(), 0)
    assert err.value.code == _engine.ERR_UNSUPPORTED and 'tensor-core' in str(err.value)
    # the frame count at the limit is accepted (tiny frames: 8192 x 130 x 200 single channel)
    n = _engine.NOALLOC_MAX_FRAMES
    d_many = torch.randint(0, 256, (n, 130, 200), dtype=torch.uint8, device='cuda')
    (want, got) = _both(eng, d_many, 1, n, 130, 200)
    _assert_same(got, want, 'n at the limit')
