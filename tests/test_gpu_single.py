"""GPU parity of the single-filter engine (through the C ABI / drop-in classes) against
the reference-generated golden vectors and the CPU oracle.  Tolerance: 1e-9 relative in
FP64 (BASELINE.json north_star); id association bit-exact."""
import numpy as np
import pytest

from conftest import GOLDEN, assert_parity, rel_err, sensitivity_floor
from oracle import ekf_oracle

pytestmark = pytest.mark.gpu
TOL = 1e-9
MODES = {"ekf_as_written": ("ekf", "as_written"), "ekf_scalar_first": ("ekf", "scalar_first"),
         "rot": ("rot", "scalar_first")}
STEP_FRAMES = (0, 1, 2, 29, 30, 31, 40, 41, 80, 119)


ENGINES = ("one_launch", "pipelined")


def make_gpu(kind, quat_mode, calibration, max_landmarks=64, engine="one_launch"):
    """engine: per-frame observe() on a small map is ONE launch of the resident kernel by default
    (asb_set_small_map_limit); "pipelined" forces the k_f1 / k_f2 / k_sweep engine that large maps use."""
    from aruco_slam_b200 import EKF, EKF_Rotations
    km, dist = calibration
    pose = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0], dtype=float)
    if kind == "ekf":
        f = EKF(pose, camera_matrix=km, dist_coeffs=dist, quat_mode=quat_mode, max_landmarks=max_landmarks)
    else:
        f = EKF_Rotations(pose, camera_matrix=km, dist_coeffs=dist, max_landmarks=max_landmarks)
    if engine == "pipelined":
        f.set_small_map_limit(0)
    return f


def order_of(landmarks):
    return [k for k, _ in sorted(landmarks.items(), key=lambda kv: kv[1])]


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("mode", list(MODES))
def test_one_step_goldens(golden_steps, calibration, mode, engine):
    g = golden_steps
    kind, qm = MODES[mode]
    f = make_gpu(kind, qm, calibration, engine=engine)
    r = 3 if kind == "ekf" else 7
    d = 3 if kind == "ekf" else 10
    for frame in STEP_FRAMES:
        key = f"{mode}/f{frame}"
        f.load_checkpoint(g[f"{key}/pre_state"], g[f"{key}/pre_P"], g[f"{key}/pre_lm"])
        ids, poses = g[f"{key}/ids"], g[f"{key}/poses"]
        f.observe(ids, poses)
        assert order_of(f.landmarks) == [int(i) for i in g[f"{key}/post_lm"]], key
        nu, blocks, _ = f.last_update()
        assert rel_err(nu, g[f"{key}/z"] - g[f"{key}/h"]) < 1e-9 * max(
            1.0, np.abs(g[f"{key}/z"]).max() / np.abs(g[f"{key}/z"] - g[f"{key}/h"]).max()), key
        # scatter the blocks into the reference's dense H
        hd = np.zeros_like(g[f"{key}/H"])
        lm = f.landmarks
        for o, idx in enumerate(ids):
            s = 10 + d * lm[int(idx)]
            hd[r * o:r * o + r, 0:10] = blocks[o][:, 0:10]
            hd[r * o:r * o + r, s:s + d] = blocks[o][:, 10:]
        assert rel_err(hd, g[f"{key}/H"]) < TOL, key
        assert_parity(f"single [{engine}] one-step vs reference golden {key}", f.state, f.uncertainty,
                      g[f"{key}/post_state"], g[f"{key}/post_P"], TOL)


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("mode", list(MODES))
def test_free_running_stream(calibration, mode, engine):
    g = np.load(GOLDEN / f"stream_{mode}.npz")
    kind, qm = MODES[mode]
    f = make_gpu(kind, qm, calibration, 16, engine=engine)
    off, cam = g["offsets"], g["cam"]
    n_frames = 40 if mode == "ekf_as_written" else len(cam)
    worst = 0.0
    for t in range(n_frames):
        a, b = off[t], off[t + 1]
        f.observe(g["ids"][a:b], g["poses"][a:b])
        if t % 25 == 0 or t == n_frames - 1:
            worst = max(worst, rel_err(f.state[:10], cam[t]))
    assert worst < TOL, worst
    if mode != "ekf_as_written":
        frames = [(g["ids"][off[t]:off[t + 1]], g["poses"][off[t]:off[t + 1]]) for t in range(n_frames)]
        floor = sensitivity_floor(kind, qm, [f_[0] for f_ in frames], [f_[1] for f_ in frames])
        assert_parity(f"single [{engine}] free-running {n_frames} frames vs reference stream golden [{mode}]", f.state,
                      f.uncertainty, g["final_state"], g["final_P"], TOL, floor=floor)
        assert order_of(f.landmarks) == [int(i) for i in g["lm"]]


@pytest.mark.parametrize("engine", ENGINES)
def test_as_written_resynchronised_every_frame(calibration, engine):
    """as-written EKF is chaotic, so every frame is compared one-step
    with (state, P) re-synchronised from the oracle (itself pinned to the reference)."""
    from aruco_slam_b200 import synth
    s = synth.make_stream(3, 10, 300, 0.8, "ekf")
    o = ekf_oracle.make("ekf", "as_written")
    f = make_gpu("ekf", "as_written", calibration, 16, engine=engine)
    worst = 0.0
    for t in range(s.n_frames):
        if t > 0:
            f.load_checkpoint(o.state, np.asarray(o.uncertainty), order_of(o.landmarks))
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
        f.observe(s.ids[t], s.poses[t])
        worst = max(worst, rel_err(f.state, o.state), rel_err(f.uncertainty, o.uncertainty))
        if t in (1, 150, s.n_frames - 1):
            assert_parity(f"single [{engine}] as-written EKF re-synchronised, frame {t}", f.state, f.uncertainty, o.state,
                          o.uncertainty, TOL)
    assert worst < TOL, worst


def test_stream_mode_equals_per_frame_calls(calibration):
    """asb_stream_upload/run (device-resident observations) == per-frame observe()."""
    from aruco_slam_b200 import synth
    s = synth.make_stream(5, 12, 60, 0.6, "ekf")
    ids, poses, off = s.packed()
    a = make_gpu("ekf", "scalar_first", calibration, 16, engine="pipelined")
    b = make_gpu("ekf", "scalar_first", calibration, 16)
    c = make_gpu("ekf", "scalar_first", calibration, 16)     # one launch per frame (small map)
    for t in range(s.n_frames):
        a.observe(s.ids[t], s.poses[t])
        c.observe(s.ids[t], s.poses[t])
    assert rel_err(a.state, c.state) < 1e-11 and rel_err(a.uncertainty, c.uncertainty) < 1e-11
    assert order_of(a.landmarks) == order_of(c.landmarks)
    b.stream_upload(ids, poses, off)
    b.stream_run(0, 30)
    b.stream_run(30, 60)
    # not bit-identical: stream mode factors S six rows per block step (k_f1<KIND, true>), per-frame
    # calls three; same arithmetic otherwise
    assert rel_err(a.state, b.state) < 1e-12
    assert rel_err(a.uncertainty, b.uncertainty) < 1e-12
    assert order_of(dict(b.get_lm_estimates())) == order_of(a.landmarks)


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("kind", ["ekf", "rot"])
def test_chunked_update_many_observations(calibration, kind, engine):
    """More rows than one Cholesky chunk holds (m > 126): chunk-sequential conditioning
    must equal the reference's single batch update.  (EKF: N = 190, the largest map the one-launch engine takes by default; EKF_Rotations: N = 250, pipelined either way.)"""
    from aruco_slam_b200 import synth
    n = 60 if kind == "ekf" else 24
    s = synth.make_stream(11, n, 6, 1.0, kind)
    o = ekf_oracle.make(kind, "scalar_first")
    f = make_gpu(kind, "scalar_first", calibration, 64, engine=engine)
    for t in range(s.n_frames):
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
        f.observe(s.ids[t], s.poses[t])
        assert_parity(f"single [{engine}] chunked update {kind} frame {t}", f.state, f.uncertainty, o.state, o.uncertainty, TOL)


@pytest.mark.parametrize("kind", ["ekf", "rot"])
def test_engines_hand_over_to_each_other(calibration, kind):
    """One filter, the per-frame engine switched back and forth (asb_set_small_map_limit): the one-launch kernel
    must find the covariance the pipelined engine left (ping-pong buffers, lazy operand) and the other way round;
    update() / predict() on their own always take the pipelined kernels."""
    from aruco_slam_b200 import synth
    s = synth.make_stream(21, 9, 60, 0.7, kind)
    o = ekf_oracle.make(kind, "scalar_first")
    f = make_gpu(kind, "scalar_first", calibration, 16)
    pattern = (4096, 0, 0, 4096, 4096, 0, 4096)
    for t in range(s.n_frames):
        f.set_small_map_limit(pattern[t % len(pattern)])
        ids, poses = [int(i) for i in s.ids[t]], s.poses[t]
        if t % 11 == 5 and all(i in o.landmarks for i in ids):
            o.predict(); o.update(ids, poses)
            f.predict(); f.update(s.ids[t], poses)
        else:
            o.observe(ids, poses)
            f.observe(s.ids[t], poses)
        if t % 7 == 0 or t == s.n_frames - 1:
            assert rel_err(f.state, o.state) < TOL, t
    assert_parity(f"single {kind}, engines alternating per frame, 60 frames", f.state, f.uncertainty, o.state,
                  o.uncertainty, TOL)
    assert order_of(f.landmarks) == order_of(o.landmarks)


def test_map_build_reduced_c3(calibration):
    """C3 at reduced size: 200 markers added 20 per frame, then updates
    of 20 known markers; full build + run parity against the oracle."""
    from aruco_slam_b200 import synth
    stream, n_build = synth.make_map_build_stream(200, 20, 8)
    o = ekf_oracle.make("ekf", "scalar_first")
    f = make_gpu("ekf", "scalar_first", calibration, 200)
    for t in range(stream.n_frames):
        o.observe([int(i) for i in stream.ids[t]], stream.poses[t])
        f.observe(stream.ids[t], stream.poses[t])
    assert f.num_landmarks == 200
    assert order_of(f.landmarks) == order_of(o.landmarks)
    assert_parity("single C3 reduced (200 markers) build + run", f.state, f.uncertainty, o.state, o.uncertainty, TOL)


@pytest.mark.parametrize("kind", ["ekf", "rot"])
def test_stream_mode_matches_oracle(calibration, kind):
    """asb_stream_run (the bench `value` path: programmatic dependent launches, k_f1 on its own SM,
    lazy operand, lower-triangle covariance) against the oracle: map build with augmentation every
    frame and frames of more than one chunk (25 > 21 markers for EKF, 12 > 8 for EKF_Rotations),
    run split over several calls."""
    from aruco_slam_b200 import synth
    if kind == "ekf":
        stream, n_build = synth.make_map_build_stream(150, 25, 10)
        frames = n_build + 10
    else:
        stream = synth.make_stream(9, 12, 25, 1.0, kind)
        frames = stream.n_frames
    o = ekf_oracle.make(kind, "scalar_first")
    for t in range(frames):
        o.observe([int(i) for i in stream.ids[t]], stream.poses[t])
    f = make_gpu(kind, "scalar_first", calibration, 150)
    ids, poses, off = stream.packed()
    f.stream_upload(ids[:off[frames]], poses[:off[frames]], off[:frames + 1])
    f.stream_run(0, 3)
    f.stream_run(3, 4)
    f.stream_run(4, frames)
    assert order_of(f.landmarks) == order_of(o.landmarks)
    assert_parity(f"single stream mode {kind}", f.state, f.uncertainty, o.state, o.uncertainty, TOL)


def test_full_size_c3_single_steps(calibration):
    """C3 at full size (2000 markers, N = 6010): build on the GPU, copy (state, P,
    landmarks) into the oracle, compare 3 single steps, plus the
    size-independent properties: P stays symmetric, diag(P) does not grow under update."""
    from aruco_slam_b200 import synth
    stream, n_build = synth.make_map_build_stream(2000, 20, 3)
    f = make_gpu("ekf", "scalar_first", calibration, 2000)
    ids, poses, off = stream.packed()
    f.stream_upload(ids, poses, off)
    f.stream_run(0, n_build)
    assert f.num_landmarks == 2000
    o = ekf_oracle.make("ekf", "scalar_first", faithful_cov=False)
    for t in range(n_build, n_build + 3):
        p_before = f.uncertainty
        o.state = f.state
        o.uncertainty = p_before.copy()
        o.landmarks = dict(f.get_lm_estimates())
        o.num_landmarks = 2000
        o.observe([int(i) for i in stream.ids[t]], stream.poses[t])
        f.stream_run(t, t + 1)
        p_after = f.uncertainty
        assert_parity(f"single C3 full size (N = 6010) one step, frame {t}", f.state, p_after, o.state,
                      o.uncertainty, TOL)
        assert np.array_equal(p_after, p_after.T)
        q = np.full(6010, 0.01)
        q[0:3], q[3:7], q[7:10] = 0.3, 0.0, 0.5
        assert np.all(np.diagonal(p_after) <= np.diagonal(p_before) + q + 1e-12)
    # the per-frame path at full size (main.run_slam's path): k_f1 on the side stream on the SM that the previous
    # frame's sweep keeps free (sweep.cuh reserve mode: CTAs on that SM retire, static ranges are claimed).  Two
    # back-to-back observe() calls so that the second front end overlaps the first sweep, then compare.
    extra, _ = synth.make_map_build_stream(2000, 20, 4, seed=99)
    o.state = f.state
    o.uncertainty = f.uncertainty
    for t in range(n_build, n_build + 4):
        f.observe(extra.ids[t], extra.poses[t])
        o.observe([int(i) for i in extra.ids[t]], extra.poses[t])
    assert_parity("single C3 full size (N = 6010), 4 per-frame observe() calls back to back", f.state, f.uncertainty,
                  o.state, o.uncertainty, TOL)


def test_api_errors_and_accessors(calibration):
    f = make_gpu("ekf", "as_written", calibration, 4)
    f.growable = False   # fixed capacity: the C ABI's ASB_ERR_CAPACITY surfaces (growth is tested in test_gpu_dropin)
    with pytest.raises(KeyError):
        f.update([5], np.zeros((1, 6)))
    f.observe([7, 9], np.array([[0, 0, 2.0, 0, 0, 0], [0.5, 0, 3.0, 0, 0, 0]]))
    assert np.allclose(f.state[3:7], [-1, 0, 0, 0])  # SURVEY.md F2 known answer
    cam, lms = f.get_poses()
    assert cam.shape == (10,) and lms.shape == (2, 3)
    assert f.get_lm_uncertainties().shape == (2, 3)
    assert dict(f.get_lm_estimates()) == {7: 0, 9: 1}
    from aruco_slam_b200._lib import AsbError
    with pytest.raises(AsbError):
        f.observe([1, 2, 3], np.zeros((3, 6)) + [0, 0, 2.0, 0, 0, 0])  # capacity 4 exceeded
    # add_marker with an explicit diagonal, predict alone
    g = make_gpu("ekf", "scalar_first", calibration, 4)
    o = ekf_oracle.make("ekf", "scalar_first")
    for flt in (g, o):
        flt.add_marker(3, np.array([0.1, 0.2, 2.0, 0, 0, 0]), np.array([0.2, 0.3, 0.4]))
        flt.predict()
        flt.update([3], np.array([[0.12, 0.2, 2.1, 0, 0, 0]]))
    assert rel_err(g.state, o.state) < TOL
    assert rel_err(g.uncertainty, o.uncertainty) < TOL


def test_missing_calibration_raises(tmp_path, monkeypatch):
    from aruco_slam_b200 import EKF
    monkeypatch.chdir(tmp_path)
    with pytest.raises(FileNotFoundError):
        EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]))


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("kind", ["ekf", "rot"])
def test_process_detections_fused_equals_separate_calls(calibration, kind, engine):
    """BASE:196-212: process_detections (IPPE + observe in one library call, poses kept on the
    device) gives exactly what estimate_pose_of_markers followed by observe gives."""
    from aruco_slam_b200 import synth
    km, dist = calibration
    a = make_gpu(kind, "as_written", calibration, engine=engine)
    b = make_gpu(kind, "as_written", calibration, engine=engine)
    rng = np.random.default_rng(5)
    corners_all = synth.make_corner_sets(400, km, dist, seed=3)
    for frame in range(12):
        k = int(rng.integers(1, 9))
        ids = rng.choice(20, size=k, replace=False).astype(np.int32)
        corners = corners_all[rng.choice(400, size=k, replace=False)]
        cam_a, lm_a, det_a = a.process_detections(corners, ids, True, frame, 0.16)
        det_b = b.estimate_pose_of_markers(corners, ids, 0.16)
        b.observe(ids, det_b)
        cam_b, lm_b = b.get_poses()
        assert np.array_equal(det_a, det_b)
        assert np.array_equal(cam_a, cam_b) and np.array_equal(lm_a, lm_b)
    assert order_of(a.landmarks) == order_of(b.landmarks)
    assert np.array_equal(a.uncertainty, b.uncertainty)


def test_stream_run_refuses_a_stale_upload(calibration):
    """asb_stream_upload bakes landmark indices against the map as it was: a map mutation in between
    (observe / add_marker / set_landmarks) must be refused, not run on stale indices."""
    from aruco_slam_b200 import synth
    from aruco_slam_b200._lib import ERR_NO_STREAM, AsbError
    s = synth.make_stream(5, 8, 10, 0.7, "ekf")
    ids, poses, off = s.packed()
    f = make_gpu("ekf", "scalar_first", calibration, 16)
    f.stream_upload(ids, poses, off)
    f.stream_run(0, 4)
    f.observe([777], np.array([[0.1, 0.2, 3.0, 0, 0, 0]]))       # a new landmark outside the stream
    with pytest.raises(AsbError) as err:
        f.stream_run(4, 10)
    assert err.value.code == ERR_NO_STREAM
    f.stream_upload(ids, poses, off)                               # a fresh upload is accepted again
    f.stream_run(0, 10)
    assert np.isfinite(f.state).all()
    # a failed upload (more landmarks than the capacity) leaves no half-built stream behind
    g = make_gpu("ekf", "scalar_first", calibration, 4)
    with pytest.raises(AsbError):
        g.stream_upload(ids, poses, off)
    with pytest.raises(AsbError) as err:
        g.stream_run(0, 1)
    assert err.value.code == ERR_NO_STREAM


def test_sweep_kernel_configurations(calibration):
    """asb_bench_sweep (measurement helper): both tile configurations of k_sweep launch and time."""
    import ctypes as C
    from aruco_slam_b200 import _lib
    f = make_gpu("ekf", "scalar_first", calibration, 300)
    f.set_landmarks(np.arange(300))
    for cfg in (0, 1):
        us = C.c_double()
        _lib.check(_lib.lib().asb_bench_sweep(f._h, 60, cfg, 0, 3, C.byref(us)))
        assert us.value > 0.0
