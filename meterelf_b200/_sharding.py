"""Frames shard by index across the GPUs of one box: one process per GPU,
contiguous index ranges, params replicated, NO data-path collective -- the only exchange
is the host-side gather of the result records (96 bytes per frame), done with
``torch.distributed`` on whatever backend the process group uses (``nccl`` on the GPU box,
``gloo`` in the CPU tests).  The reference has no counterpart (a sequential loop,
``_api.py:22``); frames are independent there too (no cross-frame state).
"""
from typing import Callable, List, Optional, Tuple

import numpy as np


def shard_bounds(n: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous ranges [k*n/G, (k+1)*n/G) for k < G; sizes differ by at most one."""
    if world < 1 or n < 0:
        raise ValueError('need world >= 1 and n >= 0')
    return [((k * n) // world, ((k + 1) * n) // world) for k in range(world)]


def gather_records(local: np.ndarray, n_total: int, dst: Optional[int] = 0) -> Optional[np.ndarray]:
    """Host-side gather of each rank's result records, in rank (= frame index) order.
    ``dst=None`` gathers on every rank.  Without an initialised process group this is the
    identity (single GPU)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        if len(local) != n_total:
            raise ValueError('single rank holds %d of %d records' % (len(local), n_total))
        return local
    (rank, world) = (dist.get_rank(), dist.get_world_size())
    bounds = shard_bounds(n_total, world)
    if len(local) != bounds[rank][1] - bounds[rank][0]:
        raise ValueError('rank %d holds %d records, its shard has %d' % (
            rank, len(local), bounds[rank][1] - bounds[rank][0]))
    item = local.dtype.itemsize
    longest = max(b - a for (a, b) in bounds)
    # equal-sized byte buffers (all_gather needs that); the NCCL backend moves device tensors
    on_gpu = dist.get_backend() == 'nccl'
    dev = torch.device('cuda', torch.cuda.current_device()) if on_gpu else torch.device('cpu')
    mine = torch.zeros(longest * item, dtype=torch.uint8)
    if len(local):
        mine[:len(local) * item] = torch.from_numpy(
            np.frombuffer(np.ascontiguousarray(local).tobytes(), np.uint8).copy())
    mine = mine.to(dev)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    if dst is not None and rank != dst:
        return None
    out = np.empty((n_total,), local.dtype)
    for (k, (a, b)) in enumerate(bounds):
        raw = parts[k].cpu().numpy()[:(b - a) * item].tobytes()
        out[a:b] = np.frombuffer(raw, local.dtype)
    return out


def read_frames_sharded(read_fn: Callable[[np.ndarray], np.ndarray], frames: np.ndarray,
                        dst: Optional[int] = 0) -> Optional[np.ndarray]:
    """Every rank passes the same ``frames`` (n, h, w, 3) array (or at least its own slice of
    it); rank k reads frames [k*n/G, (k+1)*n/G) with ``read_fn`` (``Engine.read_frames`` of
    the rank's GPU) and the records are gathered in frame order.  ``frames`` may be any
    sliceable sequence: a list of JPEG byte strings with ``Engine.read_jpeg_streams`` as
    ``read_fn`` shards the ingest path the same way."""
    import torch.distributed as dist
    n = len(frames)
    if dist.is_available() and dist.is_initialized():
        (rank, world) = (dist.get_rank(), dist.get_world_size())
    else:
        (rank, world) = (0, 1)
    (a, b) = shard_bounds(n, world)[rank]
    return gather_records(read_fn(frames[a:b]), n, dst)
