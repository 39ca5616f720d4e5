"""``get_meter_values`` keeps the reference's generator contract (``_api.py:16-33``) with batches in flight.

Host logic only: the batched reader is replaced by a stand-in, so no GPU is needed.  What the reference's loop
guarantees and the pipelined version must keep: one item per filename in input order across batch borders; a
per-file ``ImageProcessingError`` lands in ``.error``; an error the reference would have raised at file k (a
``Deferred`` slot) is raised after the items before k have been yielded; DEBUG mode stops at the first error;
an abandoned generator does not hang; at most ``PREFETCH_BATCHES`` batches are started ahead of the consumer.
"""
import threading

import pytest

from helpers import golden_set, write_params_file
from meterelf_b200 import _api, _debug
from meterelf_b200._reading import Deferred
from meterelf_b200.exceptions import DialsNotFoundError


@pytest.fixture
def params_file(tmp_path):
    return write_params_file(golden_set(1), tmp_path)


def _install(monkeypatch, reader):
    monkeypatch.setattr(_api, 'read_image_files', reader)


def test_order_across_batches_and_errors_in_their_slots(params_file, monkeypatch):
    started = []

    def reader(params, imgfs):
        started.append([f.filename for f in imgfs])
        out = []
        for f in imgfs:
            k = int(f.filename[1:])
            if k % 7 == 3:
                out.append(({}, DialsNotFoundError(f.filename, extra_info={'match val': 1.0})))
            else:
                out.append(({'value': float(k), 'a': 0.5}, None))
        return out

    _install(monkeypatch, reader)
    names = ['f%d' % k for k in range(53)]
    got = list(_api.get_meter_values(params_file, iter(names), batch_size=8))
    assert [g.filename for g in got] == names
    assert [len(b) for b in started] == [8] * 6 + [5]
    for (k, g) in enumerate(got):
        if k % 7 == 3:
            assert g.value is None and isinstance(g.error, DialsNotFoundError) and g.meter_values == {}
        else:
            assert g.value == float(k) and g.error is None and g.meter_values['a'] == 0.5


def test_deferred_error_is_raised_at_its_file_after_the_items_before_it(params_file, monkeypatch):
    def reader(params, imgfs):
        return [Deferred(ValueError('boom ' + f.filename)) if f.filename == 'f13' else ({'value': 1.0}, None)
                for f in imgfs]

    _install(monkeypatch, reader)
    gen = _api.get_meter_values(params_file, ['f%d' % k for k in range(40)], batch_size=5)
    seen = []
    with pytest.raises(ValueError, match='boom f13'):
        for item in gen:
            seen.append(item.filename)
    assert seen == ['f%d' % k for k in range(13)]


def test_debug_mode_reraises_and_reads_one_batch_at_a_time(params_file, monkeypatch):
    calls = []

    def reader(params, imgfs):
        calls.append(threading.current_thread().name)
        return [({}, DialsNotFoundError(f.filename, extra_info={'match val': 1.0})) if f.filename == 'f6'
                else ({'value': 2.0}, None) for f in imgfs]

    _install(monkeypatch, reader)
    monkeypatch.setattr(_debug, 'DEBUG', True)
    seen = []
    with pytest.raises(DialsNotFoundError):
        for item in _api.get_meter_values(params_file, ['f%d' % k for k in range(20)], batch_size=4):
            seen.append(item.filename)
    assert seen == ['f%d' % k for k in range(6)]
    assert len(calls) == 2 and not any(name.startswith('meterelf-batch') for name in calls)   # no look-ahead


def test_lookahead_is_bounded_and_an_abandoned_generator_does_not_hang(params_file, monkeypatch):
    started = []

    def reader(params, imgfs):
        started.append(imgfs[0].filename)
        return [({'value': 3.0}, None) for _ in imgfs]

    _install(monkeypatch, reader)

    def endless():
        k = 0
        while True:
            yield 'f%d' % k
            k += 1

    gen = _api.get_meter_values(params_file, endless(), batch_size=10)
    first = [next(gen).filename for _ in range(3)]
    assert first == ['f0', 'f1', 'f2']
    assert len(started) <= _api.PREFETCH_BATCHES            # nothing beyond the batches in flight was started
    done = threading.Event()

    def close():
        gen.close()
        done.set()

    t = threading.Thread(target=close)
    t.start()
    t.join(timeout=20)
    assert done.is_set()
    assert len(started) <= _api.PREFETCH_BATCHES + 1


def test_empty_input(params_file, monkeypatch):
    _install(monkeypatch, lambda params, imgfs: pytest.fail('nothing to read'))
    assert list(_api.get_meter_values(params_file, [])) == []
