"""GPU parity of the batched engine and the IPPE kernel against the CPU oracle and the
golden vectors (1e-9 relative, FP64; ids bit-exact)."""
import numpy as np
import pytest

from conftest import GOLDEN, assert_parity, rel_err, sensitivity_floor
from oracle import ekf_oracle
from oracle import ippe_oracle as io

pytestmark = pytest.mark.gpu
TOL = 1e-9


def run_oracle(kind, quat_mode, stream, n_frames):
    o = ekf_oracle.make(kind, quat_mode)
    for t in range(n_frames):
        if len(stream.ids[t]):
            o.observe([int(i) for i in stream.ids[t]], stream.poses[t])
    return o


def pack(streams, n_frames):
    ids, poses, counts = [], [], []
    for s in streams:
        for t in range(n_frames):
            ids.append(np.asarray(s.ids[t], dtype=np.int32))
            poses.append(np.asarray(s.poses[t], dtype=np.float64).reshape(-1, 6))
            counts.append(len(s.ids[t]))
    off = np.zeros(len(counts) + 1, dtype=np.int64)
    off[1:] = np.cumsum(counts)
    return np.concatenate(ids), np.concatenate(poses), off


@pytest.mark.parametrize("kind,quat_mode,n_markers,n_frames", [
    ("ekf", "scalar_first", 10, 200), ("ekf", "as_written", 10, 40), ("rot", "scalar_first", 6, 120),
    ("ekf", "scalar_first", 32, 60)])
def test_batch_matches_oracle(kind, quat_mode, n_markers, n_frames):
    from aruco_slam_b200 import FilterBatch, synth
    n_seq = 6
    streams = [synth.make_stream(20 + s, n_markers, n_frames, 0.6, kind) for s in range(n_seq)]
    # ragged input: one sequence has empty frames (no detection => no predict), one repeats an id
    streams[1].ids[3] = streams[1].ids[3][:0]
    streams[1].poses[3] = streams[1].poses[3][:0]
    streams[2].ids[5] = np.concatenate([streams[2].ids[5], streams[2].ids[5][:1]])
    streams[2].poses[5] = np.vstack([streams[2].poses[5], streams[2].poses[5][:1] + 1e-3])
    ids, poses, off = pack(streams, n_frames)
    b = FilterBatch(kind, n_seq, n_markers, n_markers + 1, quat_mode=quat_mode)
    b.upload(ids, poses, off)
    b.run(0, n_frames // 2)
    b.run(n_frames // 2, n_frames)  # state persists across runs
    b.synchronize()
    cams, status = b.cameras()
    assert not status.any()
    for s in range(n_seq):
        o = run_oracle(kind, quat_mode, streams[s], n_frames)
        order = [k for k, _ in sorted(o.landmarks.items(), key=lambda kv: kv[1])]
        assert [int(i) for i in b.landmark_ids(s)] == order
        floor = sensitivity_floor(kind, quat_mode, streams[s].ids[:n_frames], streams[s].poses[:n_frames]) \
            if s == 0 else None
        assert_parity(f"batch resident {kind}/{quat_mode} {n_markers} markers {n_frames} frames seq {s}", b.state(s),
                      b.uncertainty(s), o.state, o.uncertainty, TOL, floor=floor)
        assert rel_err(cams[s], o.state[:10]) < TOL


@pytest.mark.parametrize("kind,quat_mode,n_frames", [("ekf", "scalar_first", 40), ("rot", "scalar_first", 14)])
def test_batch_c5_shape_64_markers(kind, quat_mode, n_frames):
    """BASELINE configs[4] shape: 64 markers, p = 0.25.  EKF (N = 202) keeps P in shared memory, one
    CTA per SM; EKF_Rotations (N = 650, P 3.4 MB) runs with the covariance in HBM."""
    from aruco_slam_b200 import FilterBatch, synth
    n_seq = 3
    streams = [synth.make_stream(70 + s, 64, n_frames, 0.25, kind) for s in range(n_seq)]
    ids, poses, off = pack(streams, n_frames)
    b = FilterBatch(kind, n_seq, 64, 64, quat_mode=quat_mode)
    b.upload(ids, poses, off)
    b.run()
    _, status = b.cameras()
    assert not status.any()
    for s in range(n_seq):
        o = run_oracle(kind, quat_mode, streams[s], n_frames)
        order = [k for k, _ in sorted(o.landmarks.items(), key=lambda kv: kv[1])]
        assert [int(i) for i in b.landmark_ids(s)] == order
        assert_parity(f"batch C5 shape {kind} 64 markers {n_frames} frames seq {s}", b.state(s), b.uncertainty(s),
                      o.state, o.uncertainty, TOL)


def test_batch_full_size_c2_properties():
    """BASELINE configs[1] at full size (4096 sequences x 32 markers x 1000 frames, 64 distinct
    streams tiled as in bench.py) through the grouped upload -> run path: every sequence finite and
    status 0; tiled copies of one stream bit-identical (the engine is deterministic and sequences
    do not interact); first and last sequence within 1e-9 of the free-running oracle over all 1000
    frames (scalar-first mode: the as-written quaternion reset is chaotic)."""
    from aruco_slam_b200 import FilterBatch, synth
    n_seq, n_markers, n_frames, unique = 4096, 32, 1000, 64
    ids, poses, off = synth.make_batch(n_seq, n_markers, n_frames, 0.5, "ekf", unique=unique)
    b = FilterBatch("ekf", n_seq, n_markers, n_markers, quat_mode="scalar_first")
    b.upload(ids, poses, off)
    b.run()
    cams, status = b.cameras()
    assert not status.any() and np.isfinite(cams).all()
    for s in range(unique):
        assert np.array_equal(cams[s::unique], np.broadcast_to(cams[s], cams[s::unique].shape)), s
    for s in (0, n_seq - 1):
        stream = synth.make_stream(s % unique, n_markers, n_frames, 0.5, "ekf")
        o = run_oracle("ekf", "scalar_first", stream, n_frames)
        floor = sensitivity_floor("ekf", "scalar_first", stream.ids, stream.poses)
        assert_parity(f"batch C2 full size (4096 x 32 x 1000) seq {s}", b.state(s), b.uncertainty(s), o.state,
                      o.uncertainty, TOL, floor=floor)


@pytest.mark.parametrize("engine", ["resident", "wide"])
def test_batch_as_written_resynchronised_every_frame(engine):
    """SURVEY.md 8c(ii) for the batched engines (the headline C2 workload runs `as_written`): the as-written EKF
    is chaotic, so every frame is compared one step at a time with (state, P, id map) of every sequence
    re-synchronised from the oracle through asb_batch_set_landmarks / _set_state / _set_cov."""
    from aruco_slam_b200 import FilterBatch, synth
    n_seq, n_markers, n_frames = 3, 12, 120
    streams = [synth.make_stream(90 + s, n_markers, n_frames, 0.7, "ekf") for s in range(n_seq)]
    ids, poses, off = pack(streams, n_frames)
    b = FilterBatch("ekf", n_seq, n_markers, n_markers, quat_mode="as_written", engine=engine)
    b.upload(ids, poses, off)
    oracles = [ekf_oracle.make("ekf", "as_written") for _ in range(n_seq)]
    worst = 0.0
    for t in range(n_frames):
        if t > 0:
            for s, o in enumerate(oracles):
                order = [k for k, _ in sorted(o.landmarks.items(), key=lambda kv: kv[1])]
                b.load_checkpoint(s, o.state, np.asarray(o.uncertainty), order)
        b.run(t, t + 1)
        for s, o in enumerate(oracles):
            o.observe([int(i) for i in streams[s].ids[t]], streams[s].poses[t])
            es, ep = rel_err(b.state(s), o.state), rel_err(b.uncertainty(s), o.uncertainty)
            worst = max(worst, es, ep)
            if t in (1, n_frames // 2, n_frames - 1) and s == 0:
                assert_parity(f"batch {engine} as-written EKF re-synchronised, frame {t}", b.state(s), b.uncertainty(s),
                              o.state, o.uncertainty, TOL)
    assert worst < TOL, worst


def test_batch_equals_single_filter(calibration):
    """The two engines are independent implementations of the same update."""
    from aruco_slam_b200 import EKF, FilterBatch, synth
    km, dist = calibration
    s = synth.make_stream(42, 12, 80, 0.7, "ekf")
    ids, poses, off = pack([s], 80)
    b = FilterBatch("ekf", 1, 12, 12, quat_mode="scalar_first")
    b.upload(ids, poses, off)
    b.run()
    f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist,
            quat_mode="scalar_first", max_landmarks=12)
    for t in range(80):
        f.observe(s.ids[t], s.poses[t])
    assert rel_err(b.state(0), f.state) < 1e-10
    assert rel_err(b.uncertainty(0), f.uncertainty) < 1e-10


def test_batch_capacity_flag():
    from aruco_slam_b200 import FilterBatch, synth
    s = synth.make_stream(1, 10, 5, 1.0, "ekf")
    ids, poses, off = pack([s], 5)
    b = FilterBatch("ekf", 1, 4, 16)  # 10 markers > max_landmarks 4
    b.upload(ids, poses, off)
    b.run()
    _, status = b.cameras()
    assert status[0] == -4


def _check_pose(out, ref):
    et = np.abs(out[:, :3] - ref[:, :3]).max(1) / np.abs(ref[:, :3]).max(1)
    er = np.abs(io.rodrigues_matrix(out[:, 3:]) - io.rodrigues_matrix(ref[:, 3:])) \
        .reshape(len(out), -1).max(1)
    return et, er


def test_ippe_golden_cv2():
    from aruco_slam_b200 import estimate_pose_of_markers
    g = np.load(GOLDEN / "ippe_golden.npz")
    out = estimate_pose_of_markers(g["corners"], g["camera_matrix"], g["dist_coeffs"],
                                   float(g["marker_size"]))
    et, er = _check_pose(out, g["pose"])
    assert et.max() < TOL and er.max() < TOL


def test_ippe_both_solutions_golden_cv2():
    """cv2.solvePnPGeneric(..., SOLVEPNP_IPPE_SQUARE): both solutions, better one first.  The second
    solution of a near-frontal marker is ill-conditioned (same bound as the oracle's own test)."""
    from aruco_slam_b200.ippe import estimate_both_poses, estimate_pose_of_markers
    g = np.load(GOLDEN / "ippe_golden.npz")
    args = (g["corners"], g["camera_matrix"], g["dist_coeffs"], float(g["marker_size"]))
    both, errs = estimate_both_poses(*args)
    assert both.shape == (2000, 2, 6) and errs.shape == (2000, 2)
    assert np.array_equal(both[:, 0], estimate_pose_of_markers(*args))
    assert (errs[:, 0] <= errs[:, 1]).all()
    for j in range(2):
        et, er = _check_pose(both[:, j], g["both"][:, j])
        assert np.median(et) < 1e-12 and et.max() < 1e-7 and er.max() < 1e-7, j


def test_ippe_large_batch_vs_oracle(calibration):
    from aruco_slam_b200 import estimate_pose_of_markers, synth
    km, dist = calibration
    corners = synth.make_corner_sets(200_000, km, dist, seed=7)
    out = estimate_pose_of_markers(corners, km, dist, 0.16)
    ref = io.estimate_pose_of_markers(corners, km, dist, 0.16)
    et, er = _check_pose(out, ref)
    # solution ordering is decided by a float32 comparison; allow a vanishing fraction of
    # near-ties to flip (none seen on cv2's own 2000 golden sets)
    bad = (et > TOL) | (er > TOL)
    assert bad.mean() < 1e-4, bad.sum()
    assert np.isfinite(out).all()
    # round trip: reprojecting the recovered pose reproduces the corners to pixel noise
    h = 0.08
    obj = np.array([[-h, h, 0], [h, h, 0], [h, -h, 0], [-h, -h, 0]])
    uv, _ = synth.project_points(obj, out[:2000, 3:], out[:2000, :3], km, dist)
    err = np.abs(uv - corners[:2000])
    assert np.median(err) < 0.5 and err.max() < 8.0  # pixel noise is N(0, 0.3^2)


def test_ippe_sliced_host_path_equals_single_shot(calibration):
    """Batches of >= 262144 sets go through the three-stream slice pipeline of ippe_host (32768-set slices, a ring of
    64 events): 2.3 M sets = 71 slices, so the ring wraps and the last slice is ragged.  Bit-exact against the same
    sets sent in single-shot pieces of 200 000, for the one-solution and the two-solution entry points."""
    from aruco_slam_b200 import estimate_pose_of_markers, synth
    from aruco_slam_b200.ippe import estimate_both_poses
    km, dist = calibration
    base = synth.make_corner_sets(230_000, km, dist, seed=11)
    corners = np.concatenate([base + np.float32(0.25 * i) for i in range(10)])
    n = len(corners)
    assert n == 2_300_000 and n % 32768 != 0
    out = estimate_pose_of_markers(corners, km, dist, 0.16)
    piece = 200_000
    ref = np.concatenate([estimate_pose_of_markers(corners[o:o + piece], km, dist, 0.16) for o in range(0, n, piece)])
    assert np.array_equal(out, ref)
    m = 600_000
    both, errs = estimate_both_poses(corners[:m], km, dist, 0.16)
    rb, re = zip(*[estimate_both_poses(corners[o:o + piece], km, dist, 0.16) for o in range(0, m, piece)])
    assert np.array_equal(both, np.concatenate(rb)) and np.array_equal(errs, np.concatenate(re))
    assert np.array_equal(both[:, 0], out[:m])


def test_ippe_accepts_detector_layout(calibration):
    """detectMarkers returns a tuple of (1,4,2) float32 arrays (base_filter.py:194)."""
    from aruco_slam_b200 import estimate_pose_of_markers, synth
    km, dist = calibration
    corners = synth.make_corner_sets(5, km, dist, seed=3)
    as_tuple = tuple(c.reshape(1, 4, 2) for c in corners)
    a = estimate_pose_of_markers(as_tuple, km, dist, 0.16)
    b = estimate_pose_of_markers(corners, km, dist, 0.16)
    assert a.shape == (5, 6) and np.array_equal(a, b)
    assert estimate_pose_of_markers(np.zeros((0, 4, 2), np.float32), km, dist).shape == (0, 6)
