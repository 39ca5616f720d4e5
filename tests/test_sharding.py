"""Multi-GPU host logic on CPU: world_size-2 gloo process group, frames sharded by index,
host-side gather of the result records (no data-path collective).  The per-shard reader in
these tests is the oracle (tests may use it as the stand-in for ``Engine.read_frames``)."""
import os
import socket
import sys

import numpy as np
import pytest

from meterelf_b200 import _engine, _sharding


def test_shard_bounds_cover_every_index_once():
    for n in (0, 1, 2, 7, 8, 4096, 32768, 32769):
        for world in (1, 2, 3, 4, 8):
            b = _sharding.shard_bounds(n, world)
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[k][1] == b[k + 1][0] for k in range(world - 1))
            sizes = [e - s for (s, e) in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        _sharding.shard_bounds(4, 0)


def test_single_rank_gather_is_identity():
    rec = np.zeros((5,), _engine.RESULT_DTYPE)
    rec['value'] = np.arange(5)
    assert _sharding.gather_records(rec, 5) is rec
    with pytest.raises(ValueError):
        _sharding.gather_records(rec, 6)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_frames, out_dir):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle
    from helpers import golden_set
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        gs = golden_set(2)
        crops = gs.subset()['crops'][:n_frames]
        O = gs.oracle_params()

        def read_fn(frames):
            res = oracle.read_batch(O, frames, n_threads=1)
            out = np.zeros((len(frames),), _engine.RESULT_DTYPE)
            for k in out.dtype.names:
                out[k] = res[k]
            return out
        got = _sharding.read_frames_sharded(read_fn, crops, dst=None)
        np.save(os.path.join(out_dir, 'rank%d.npy' % rank), got)
        only0 = _sharding.read_frames_sharded(read_fn, crops, dst=0)
        assert (only0 is None) == (rank != 0)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_frames', [7, 2, 1])
def test_two_rank_gloo_shard_and_gather(tmp_path, n_frames):
    import torch.multiprocessing as mp
    import oracle
    from helpers import golden_set
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_frames, str(tmp_path)), nprocs=world, join=True)
    gs = golden_set(2)
    crops = gs.subset()['crops'][:n_frames]
    want = oracle.read_batch(gs.oracle_params(), crops, n_threads=1)
    for rank in range(world):
        got = np.load(os.path.join(str(tmp_path), 'rank%d.npy' % rank))
        assert len(got) == n_frames
        for k in ('status', 'match_x', 'match_y', 'failed_dial', 'unreadable_mask'):
            assert np.array_equal(got[k], want[k]), k
        assert got['dial_pos'].tobytes() == np.asarray(want['dial_pos']).tobytes()
        assert got['value'].tobytes() == np.asarray(want['value']).tobytes()


def _stream_worker(rank, world, port, n_streams, out_dir):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle
    from helpers import GOLDEN, golden_set
    from test_jpeg_oracle import load_streams
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        gs = golden_set(1)
        streams = [s[1] for s in load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
                   if s[4] == 1 and s[2] == (640, 480)][:n_streams]
        (x0, y0, x1, y1) = gs.meter_rect

        def read_fn(part):          # stand-in for Engine.read_jpeg_streams: decode + read with the oracle
            out = np.zeros((len(part),), _engine.RESULT_DTYPE)
            if len(part):
                crops = np.stack([oracle.decode_jpeg(d)[y0:y1, x0:x1] for d in part])
                res = oracle.read_batch(gs.oracle_params(), np.ascontiguousarray(crops), n_threads=1)
                for k in out.dtype.names:
                    out[k] = res[k]
            return out
        got = _sharding.read_frames_sharded(read_fn, streams, dst=None)
        np.save(os.path.join(out_dir, 'srank%d.npy' % rank), got)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_shards_jpeg_streams(tmp_path):
    """The ingest path shards like the frame path: a list of JPEG byte strings split by index."""
    import torch.multiprocessing as mp
    from helpers import GOLDEN, format_result_line, golden_set
    from test_jpeg_oracle import load_streams
    (world, n) = (2, 3)          # the committed subset holds three portrait files of set 1
    mp.spawn(_stream_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    gs = golden_set(1)
    names = [s[0] for s in load_streams(os.path.join(GOLDEN, 'jpeg_subset.npz'))
             if s[4] == 1 and s[2] == (640, 480)][:n]
    a = np.load(os.path.join(str(tmp_path), 'srank0.npy'))
    b = np.load(os.path.join(str(tmp_path), 'srank1.npy'))
    assert a.tobytes() == b.tobytes() and len(a) == n
    for (rec, name) in zip(a, names):
        exp = gs.expected[name]
        if exp['status'] == 0:
            assert format_result_line(rec, gs.names) == exp['line'], name
