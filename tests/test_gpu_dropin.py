"""GPU: the rows of SURVEY.md 8(f) — what `main/run_slam.py` / `main/run_offline.py` need from the drop-in beyond
observe(): byte-compatible `save_map` / trajectory text, the pose history (`get_cam_estimate`) behind
run_offline's second pass with a no-op `batch_optimize`, the orientation gate `correct_poses`, and unbounded
landmark growth.  Numbers against the CPU oracle at 1e-9; ids and file layout exact."""
import re

import numpy as np
import pytest

from conftest import GOLDEN, rel_err
from oracle import ekf_oracle, gate_oracle

pytestmark = pytest.mark.gpu
TOL = 1e-9
POSE0 = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])


def _gpu(kind, calibration, **kw):
    from aruco_slam_b200 import EKF, EKF_Rotations
    km, dist = calibration
    if kind == "ekf":
        return EKF(POSE0, camera_matrix=km, dist_coeffs=dist, **kw)
    return EKF_Rotations(POSE0, camera_matrix=km, dist_coeffs=dist, **kw)


def _reference_writer(path, marker_poses, index_to_id, uncertainties):
    """The reference's save_map body (base_filter.py:228-247) applied to the oracle's numbers."""
    with open(path, "w", encoding="utf-8") as file:
        file.write("# landmark_id\n# x y z\n# uncertainty\n\n")
        for i, pose in enumerate(marker_poses):
            file.write(f"{index_to_id[i]}\n")
            file.write(f"{', '.join(map(str, pose))}\n")
            file.write(f"{', '.join(map(str, uncertainties[i, :len(pose)]))}\n")
            file.write("\n")


@pytest.mark.parametrize("kind", ["ekf", "rot"])
def test_save_map_matches_reference_format_and_oracle_numbers(tmp_path, calibration, kind):
    from aruco_slam_b200 import synth
    d = 3 if kind == "ekf" else 10
    s = synth.make_stream(21, 9, 50, 0.7, kind)
    f = _gpu(kind, calibration, quat_mode="scalar_first") if kind == "ekf" else _gpu(kind, calibration)
    o = ekf_oracle.make(kind, "scalar_first")
    for t in range(s.n_frames):
        f.observe(s.ids[t], s.poses[t])
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
    f.save_map(tmp_path / "gpu.txt")
    unc = np.diagonal(np.asarray(o.uncertainty))[10:].reshape(-1, d)
    _reference_writer(tmp_path / "ref.txt", o.state[10:].reshape(-1, d), {v: k for k, v in o.landmarks.items()}, unc)
    got = (tmp_path / "gpu.txt").read_text().splitlines()
    ref = (tmp_path / "ref.txt").read_text().splitlines()
    assert len(got) == len(ref) == 4 + 4 * len(o.landmarks)
    assert got[:4] == ref[:4] == ["# landmark_id", "# x y z", "# uncertainty", ""]
    # the shipped sample of the reference (outputs/map.txt) has the same layout
    sample = (GOLDEN / "map_sample.txt").read_text().splitlines()
    assert sample[:4] == got[:4] and re.fullmatch(r"-?\d+", sample[4]) and sample[5].count(", ") == 2
    for i in range(4, len(ref), 4):
        assert got[i] == ref[i]                       # landmark id line: byte-exact
        assert got[i + 3] == ref[i + 3] == ""
        for line_g, line_r in ((got[i + 1], ref[i + 1]), (got[i + 2], ref[i + 2])):
            a, b = line_g.split(", "), line_r.split(", ")     # separator and token count byte-exact
            assert len(a) == len(b) == d
            assert rel_err(np.array(a, float), np.array(b, float)) < TOL
    # load_map (dead code in the reference, BASE:249-272) round trip: ids, positions, diagonals
    g = _gpu(kind, calibration, quat_mode="scalar_first") if kind == "ekf" else _gpu(kind, calibration)
    g.load_map(tmp_path / "gpu.txt")
    assert [k for k, _ in sorted(dict(g.get_lm_estimates()).items(), key=lambda kv: kv[1])] == \
        [k for k, _ in sorted(o.landmarks.items(), key=lambda kv: kv[1])]
    assert g.get_lm_uncertainties().shape == (len(o.landmarks), d)
    assert np.allclose(g.get_lm_uncertainties(), unc.astype(np.float32), rtol=1e-6)


def test_trajectory_writer_line_format(tmp_path, calibration):
    """outputs/trajectory_writer.py:29-40 and the sample outputs/trajectory.txt:1-3."""
    from aruco_slam_b200 import TrajectoryWriter, synth
    s = synth.make_stream(4, 6, 5, 1.0, "ekf")
    f = _gpu("ekf", calibration)
    poses = []
    with TrajectoryWriter(str(tmp_path / "traj.txt")) as w:
        for t in range(s.n_frames):
            cam, _, _ = f.process_detections(None, None) if t == 0 else (None, None, None)
            f.observe(s.ids[t], s.poses[t])
            cam, _ = f.get_poses()
            w.write(1000.0 * t / 30.0, cam)
            poses.append(cam)
    lines = (tmp_path / "traj.txt").read_text().splitlines()
    sample = (GOLDEN / "trajectory_sample.txt").read_text().splitlines()
    num = r"-?\d+\.\d+(e-?\d+)?"
    pat = re.compile(rf"\d+\.\d{{4}}( {num}){{7}}")
    assert all(pat.fullmatch(x) for x in sample), sample
    assert len(lines) == s.n_frames and all(pat.fullmatch(x) for x in lines), lines
    for t, line in enumerate(lines):
        tok = line.split(" ")
        assert tok[0] == f"{t / 30.0:.4f}"
        assert [float(x) for x in tok[1:]] == [float(v) for v in poses[t][:7]]      # str() round-trips float64
        assert tok[1:] == [str(v) for v in poses[t][:7]]                            # the reference's formatting
    assert sample[0] == "0.0000 0.0 0.0 0.0 1.0 0.0 0.0 0.0"


def test_run_offline_two_pass_loop(calibration):
    """main/run_offline.py:86-122 shape with an EKF in place of the FactorGraph: pass 1 filters and records,
    batch_optimize() is a no-op, pass 2 (should_filter=False, iteration=i) replays get_cam_estimate(i)
    (BASE:205-210, 369-381) without touching the filter."""
    from aruco_slam_b200 import synth
    km, dist = calibration
    f = _gpu("ekf", calibration, quat_mode="scalar_first")
    rng = np.random.default_rng(9)
    corners_all = synth.make_corner_sets(300, km, dist, seed=5)
    frames = []
    for i in range(15):
        k = int(rng.integers(0, 7))          # k = 0: a frame without detections (ids is None upstream)
        ids = rng.choice(12, size=k, replace=False).astype(np.int32) if k else None
        frames.append((corners_all[rng.choice(300, size=k, replace=False)] if k else (), ids))
    pass1 = []
    for i, (corners, ids) in enumerate(frames):
        cam, lms, det = f.process_detections(corners, ids)
        pass1.append(cam.copy())
        assert det.shape == ((len(ids), 6) if ids is not None else (0,))
    state_before, cov_before = f.state, f.uncertainty
    assert f.batch_optimize() is None
    for i, (corners, ids) in enumerate(frames):
        cam, lms, det = f.process_detections(corners, ids, should_filter=False, iteration=i)
        assert np.array_equal(cam, pass1[i])
        assert np.array_equal(cam, f.get_cam_estimate(i))
        assert lms.shape == (f.num_landmarks, 3)
    assert np.array_equal(f.state, state_before) and np.array_equal(f.uncertainty, cov_before)


def test_correct_poses_gate_matches_reference_restatement(calibration):
    """BASE:274-325 on the device (asb_correct_poses) against oracle/gate_oracle.py: consistent observations
    pass, an orientation flipped by more than QUAT_THRESHOLD is dropped, unknown ids are dropped; with the second
    IPPE solution supplied, a rejected first solution is replaced when the second passes."""
    from aruco_slam_b200 import synth
    s = synth.make_stream(31, 8, 30, 1.0, "rot")
    f = _gpu("rot", calibration)
    o = ekf_oracle.make("rot", "scalar_first")
    for t in range(s.n_frames - 1):
        f.observe(s.ids[t], s.poses[t])
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
    ids = np.concatenate([s.ids[-1], [9001, 9002]]).astype(np.int32)
    good = np.vstack([s.poses[-1], s.poses[-1][:2]])
    bad = good.copy()
    bad[:, 3:] += np.array([2.0, -1.5, 1.0])          # well beyond 50 degrees
    mixed = good.copy()
    mixed[::2] = bad[::2]
    rng = np.random.default_rng(3)
    border = good.copy()
    border[:, 3:] += rng.normal(0.0, 0.5, size=(len(ids), 3))   # a spread of angles around the threshold
    for poses, alt in ((good, None), (bad, None), (mixed, None), (mixed, good), (border, None), (border, good)):
        ref_ids, ref_poses, _ = gate_oracle.correct_poses(o.state, o.landmarks, ids, poses, alt)
        got_ids, got_poses = f.correct_poses(ids, poses, alt)
        assert got_ids == ref_ids
        assert len(got_poses) == len(ref_poses)
        assert all(np.array_equal(a, b) for a, b in zip(got_poses, ref_poses))
    assert f.correct_poses(ids, good)[0] == [int(i) for i in s.ids[-1]]          # unknown ids dropped
    assert f.correct_poses(ids, bad)[0] == []
    # the gate only reads the state
    assert rel_err(f.state, o.state) < TOL


def test_landmarks_grow_past_the_initial_capacity(calibration):
    """The reference grows without bound (EKF:279-290); the drop-in starts at max_landmarks and moves to a larger
    handle on overflow, copying state / covariance / id map."""
    from aruco_slam_b200 import synth
    for kind in ("ekf", "rot"):
        s = synth.make_stream(13, 21, 12, 0.8, kind)
        f = _gpu(kind, calibration, max_landmarks=4, quat_mode="scalar_first") if kind == "ekf" else \
            _gpu(kind, calibration, max_landmarks=4)
        o = ekf_oracle.make(kind, "scalar_first")
        for t in range(s.n_frames):
            f.observe(s.ids[t], s.poses[t])
            o.observe([int(i) for i in s.ids[t]], s.poses[t])
        assert f.num_landmarks == 21 and f.max_landmarks >= 21
        assert dict(f.get_lm_estimates()) == o.landmarks
        assert rel_err(f.state, o.state) < TOL
        assert rel_err(f.uncertainty, o.uncertainty) < TOL


def test_moving_average_motion_model_is_opt_in(calibration):
    """README.md:55-56 describes a moving-average motion model that the reference code does not
    contain (F4).  Off by default (the goldens pin that); switched on, every observe() first moves the camera
    position by (X_k-1 - X_k-n) / n - checked against the same step applied to the oracle's state by hand."""
    from aruco_slam_b200 import synth
    from aruco_slam_b200._lib import AsbError
    window = 4
    s = synth.make_stream(17, 8, 40, 0.8, "ekf")
    f = _gpu("ekf", calibration, quat_mode="scalar_first")
    f.set_motion_model(window)
    o = ekf_oracle.make("ekf", "scalar_first")
    hist = []
    for t in range(s.n_frames):
        cur = o.state[:3].copy()
        hist.append(cur)
        held = hist[-window:]
        if len(held) >= 2:
            o.state[:3] = cur + (cur - held[0]) / len(held)
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
        f.observe(s.ids[t], s.poses[t])
        if t in (1, 5, s.n_frames - 1):
            assert rel_err(f.state, o.state) < TOL, t
    assert rel_err(f.uncertainty, o.uncertainty) < TOL
    ids, poses, off = s.packed()
    f.stream_upload(ids, poses, off)
    with pytest.raises(AsbError):
        f.stream_run(0, 1)               # per-frame only
    f.set_motion_model(0)                # off again: the reference's predict
