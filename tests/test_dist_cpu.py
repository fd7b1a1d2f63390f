"""N > 1 host logic on CPU: world_size-2 gloo processes exercise the sharding and the final
gather that the multi-GPU bench path uses (no GPU, no compute calls)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from aruco_slam_b200 import dist as asb_dist  # noqa: E402
from aruco_slam_b200 import synth  # noqa: E402


@pytest.mark.parametrize("n,world", [(10, 1), (10, 2), (7, 4), (3, 8), (4096, 8), (0, 2)])
def test_shard_range_partitions(n, world):
    spans = [asb_dist.shard_range(n, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for (a, b), (c, d) in zip(spans, spans[1:]):
        assert b == c and a <= b and c <= d
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1


def test_shard_range_rejects_bad_rank():
    with pytest.raises(ValueError):
        asb_dist.shard_range(4, 2, 2)


def test_shard_offsets_reassemble_stream():
    ids, poses, off = synth.make_batch(6, 5, 7, 0.6, "ekf")
    got_ids, got_poses = [], []
    for r in range(3):
        lo, hi = asb_dist.shard_range(6, r, 3)
        a, b, loc = asb_dist.shard_offsets(off, 7, lo, hi)
        assert loc[0] == 0 and loc.size == (hi - lo) * 7 + 1 and loc[-1] == b - a
        got_ids.append(ids[a:b])
        got_poses.append(poses[a:b])
    assert np.array_equal(np.concatenate(got_ids), ids)
    assert np.array_equal(np.concatenate(got_poses), poses)


def _worker(rank, world, port, n_total, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = asb_dist.shard_range(n_total, rank, world)
    local = np.arange(lo, hi, dtype=np.float64)[:, None] * np.ones((1, 10)) + 0.25 * rank
    out = asb_dist.gather_results(local, n_total, dist)
    q.put((rank, None if out is None else out.tolist()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
@pytest.mark.parametrize("n_total", [7, 8])
def test_gather_results_gloo_world2(n_total):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=90) for _ in procs)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert res[1] is None
    got = np.array(res[0])
    assert got.shape == (n_total, 10)
    lo1, _ = asb_dist.shard_range(n_total, 1, 2)
    want = np.arange(n_total, dtype=np.float64)[:, None] * np.ones((1, 10))
    want[lo1:] += 0.25
    assert np.array_equal(got, want)


def test_gather_results_single_process_identity():
    a = np.arange(30.0).reshape(3, 10)
    assert asb_dist.gather_results(a, 3) is a or np.array_equal(asb_dist.gather_results(a, 3), a)


# ---- row-sharded filter (filter_shard.cu): the partition and the row-block gather, on CPU -------
@pytest.mark.parametrize("kind,max_lm,world", [("ekf", 2000, 1), ("ekf", 2000, 2), ("ekf", 2000, 8),
                                               ("rot", 64, 4), ("ekf", 1, 8), ("rot", 2000, 3)])
def test_shard_partition_covers_state(kind, max_lm, world):
    from aruco_slam_b200 import sharded
    d = 3 if kind == "ekf" else 10
    spans = [sharded.partition(kind, max_lm, world, r) for r in range(world)]
    ld = spans[0][2]
    assert spans[0][0] == 0 and spans[-1][1] == ld and ld >= 10 + d * max_lm
    for (a, b, l1), (c, e, l2) in zip(spans, spans[1:]):
        assert b == c and l1 == l2
    sizes = {b - a for a, b, _ in spans}
    assert len(sizes) == 1 and sizes.pop() % 64 == 0   # whole 64-column panel blocks, equal on every rank


def test_shard_partition_rejects_bad_rank():
    from aruco_slam_b200 import _lib, sharded
    with pytest.raises(_lib.AsbError):
        sharded.partition("ekf", 10, 2, 2)
