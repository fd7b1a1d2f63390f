"""Multi-GPU plumbing for the parts of the path that shard.

Independent sequences (BASELINE configs C2 / C5) and independent corner sets (C4) are
embarrassingly parallel: one process per GPU, contiguous blocks of sequences per rank, NO
collective on the data path, one final gather of the per-sequence results (camera pose,
status) to rank 0.  ``torch.distributed`` is plumbing only (NCCL over NVLink on the GPU box,
gloo in the CPU tests); the filters themselves never see it.
"""
from __future__ import annotations

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of ``n_total`` items owned by ``rank``; the first
    ``n_total % world`` ranks get one extra item (ragged totals are allowed, a rank may be
    empty when n_total < world)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} of world {world}")
    base, extra = divmod(int(n_total), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_offsets(offsets: np.ndarray, n_frames: int, lo: int, hi: int):
    """Slice a packed observation stream (offsets of n_seq * n_frames + 1 entries, see
    ``FilterBatch.upload``) down to sequences [lo, hi): returns (row_lo, row_hi, local_offsets)."""
    off = np.asarray(offsets, dtype=np.int64)
    a, b = lo * n_frames, hi * n_frames
    row_lo, row_hi = int(off[a]), int(off[b])
    return row_lo, row_hi, off[a:b + 1] - row_lo


def gather_results(local: np.ndarray, n_total: int, dist=None, device=None):
    """Final gather of per-sequence rows (e.g. the (n_local, 10) camera poses) to rank 0.
    Ranks may hold different row counts (``shard_range``).  Returns the (n_total, ...) array on
    rank 0 and None elsewhere; with ``dist is None`` (single process) it is the identity."""
    local = np.ascontiguousarray(local)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        assert local.shape[0] == n_total
        return local
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    counts = [shard_range(n_total, r, world) for r in range(world)]
    width = int(np.prod(local.shape[1:], dtype=np.int64)) if local.ndim > 1 else 1
    longest = max(hi - lo for lo, hi in counts)
    pad = torch.zeros((longest, width), dtype=torch.from_numpy(local.reshape(local.shape[0], -1)).dtype)
    pad[:local.shape[0]] = torch.from_numpy(local.reshape(local.shape[0], -1))
    if device is not None:
        pad = pad.to(device)
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad)          # NCCL has no ragged gather: pad to the longest shard
    if rank != 0:
        return None
    rows = [bufs[r][:hi - lo].cpu().numpy() for r, (lo, hi) in enumerate(counts)]
    return np.concatenate(rows, 0).reshape((n_total,) + local.shape[1:])
