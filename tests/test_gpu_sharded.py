"""GPU parity of the row-sharded filter (filter_shard.cu): same stream through ShardedFilter and
through the CPU oracle / the single-GPU engine.  World size 1 runs on any GPU box (all kernels
of the sharded path, collectives degenerate); world size 2 spawns one process per GPU and is
skipped on a one-GPU box.  Tolerance 1e-9 relative (FP64), id order bit-exact."""
import os
import socket

import numpy as np
import pytest

from conftest import assert_parity, rel_err
from oracle import ekf_oracle

pytestmark = pytest.mark.gpu
TOL = 1e-9
POSE0 = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0], dtype=float)


def _order(landmarks):
    return [k for k, _ in sorted(landmarks.items(), key=lambda kv: kv[1])]


@pytest.mark.parametrize("kind,quat_mode,n_markers,n_frames", [
    ("ekf", "scalar_first", 30, 40), ("ekf", "as_written", 12, 25), ("rot", "scalar_first", 14, 30)])
def test_world1_matches_oracle(kind, quat_mode, n_markers, n_frames):
    from aruco_slam_b200 import synth
    from aruco_slam_b200.sharded import ShardedFilter
    s = synth.make_stream(3, n_markers, n_frames, 0.7, kind)
    f = ShardedFilter(kind, POSE0, n_markers, quat_mode=quat_mode)
    o = ekf_oracle.make(kind, quat_mode)
    for t in range(s.n_frames):
        f.observe(s.ids[t], s.poses[t])
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
    assert _order(f.landmarks) == _order(o.landmarks)
    assert_parity("row-sharded world 1 vs oracle", f.state, f.uncertainty, o.state, o.uncertainty, TOL)


def test_world1_map_build_equals_single_gpu_engine(calibration):
    """Several row blocks, augmentation every frame, chunked updates (25 markers > kc_max = 21)."""
    from aruco_slam_b200 import EKF, synth
    from aruco_slam_b200.sharded import ShardedFilter
    km, dist = calibration
    stream, n_build = synth.make_map_build_stream(200, 25, 12)
    ids, poses, off = stream.packed()
    a = ShardedFilter("ekf", POSE0, 200)
    a.stream_upload(ids, poses, off)
    a.stream_run(0, n_build + 12)
    b = EKF(POSE0, camera_matrix=km, dist_coeffs=dist, quat_mode="as_written", max_landmarks=200)
    b.stream_upload(ids, poses, off)
    b.stream_run(0, n_build + 12)
    assert _order(a.landmarks) == _order(b.landmarks)
    assert rel_err(a.state, b.state) < TOL
    assert rel_err(a.uncertainty, b.uncertainty) < TOL


def test_world1_full_size_equals_single_gpu_engine(calibration):
    """C3 at full size (2000 markers, N = 6010) on the row-sharded engine at world 1: the rectangular sweep then has
    more tiles than CTAs and takes its contiguous-range + dynamic-tail schedule (the path the `c3_sharded` bench
    leg runs; small maps never reach it)."""
    from aruco_slam_b200 import EKF, synth
    from aruco_slam_b200.sharded import ShardedFilter
    km, dist = calibration
    stream, n_build = synth.make_map_build_stream(2000, 20, 6)
    ids, poses, off = stream.packed()
    a = ShardedFilter("ekf", POSE0, 2000, quat_mode="scalar_first")
    a.stream_upload(ids, poses, off)
    a.stream_run(0, n_build + 6)
    b = EKF(POSE0, camera_matrix=km, dist_coeffs=dist, quat_mode="scalar_first", max_landmarks=2000)
    b.stream_upload(ids, poses, off)
    b.stream_run(0, n_build + 6)
    assert_parity("row-sharded world 1 at full size (N = 6010) vs the single-GPU engine", a.state, a.uncertainty,
                  b.state, b.uncertainty, TOL)
    a.close()


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from aruco_slam_b200 import synth
    from aruco_slam_b200.sharded import ShardedFilter
    out = {}
    for kind, qm, n_markers, per_frame, extra in (("ekf", "as_written", 120, 24, 10), ("rot", "scalar_first", 40, 8, 10)):
        if kind == "ekf":
            stream, n_build = synth.make_map_build_stream(n_markers, per_frame, extra)
            frames = n_build + extra
        else:
            stream = synth.make_stream(5, n_markers, 30, 0.4, kind)
            frames = stream.n_frames
        f = ShardedFilter(kind, POSE0, n_markers, dist=dist, rank=rank, world=world, device=rank, quat_mode=qm)
        if kind == "ekf":
            # stream mode: the front of every chunk overlaps the previous chunk's sweep (look-ahead across frames)
            ids, poses, off = stream.packed()
            f.stream_upload(ids[:off[frames]], poses[:off[frames]], off[:frames + 1])
            f.stream_run(0, frames // 2)
            f.stream_run(frames // 2, frames)
        else:
            for t in range(frames):
                f.observe(stream.ids[t], stream.poses[t])
        out[kind] = (f.state, f.uncertainty, _order(f.landmarks), (f.row0, f.row1))
        f.close()
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_world2_matches_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    from aruco_slam_b200 import synth
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=500) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # the oracle on the same streams
    stream, n_build = synth.make_map_build_stream(120, 24, 10)
    o = ekf_oracle.make("ekf", "as_written")
    for t in range(n_build + 10):
        o.observe([int(i) for i in stream.ids[t]], stream.poses[t])
    s2 = synth.make_stream(5, 40, 30, 0.4, "rot")
    o2 = ekf_oracle.make("rot", "scalar_first")
    for t in range(s2.n_frames):
        o2.observe([int(i) for i in s2.ids[t]], s2.poses[t])
    for rank in (0, 1):
        for kind, orc in (("ekf", o), ("rot", o2)):
            state, cov, order, rows = res[rank][kind]
            assert order == _order(orc.landmarks), (rank, kind)
            assert rel_err(state, orc.state) < TOL, (rank, kind)
            assert rel_err(cov, np.asarray(orc.uncertainty)) < TOL, (rank, kind)
    assert res[0]["ekf"][3][1] == res[1]["ekf"][3][0]  # contiguous row ranges
