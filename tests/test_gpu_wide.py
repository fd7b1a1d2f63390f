"""GPU parity of the wide batched engine (frame-synchronous kernels, covariance tiled in HBM:
wide_kernels.cuh) against the CPU oracle: 1e-9 relative in FP64, id order bit-exact."""
import numpy as np
import pytest

from conftest import assert_parity, rel_err, sensitivity_floor
from test_gpu_batch_ippe import pack, run_oracle

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _check(b, streams, kind, quat_mode, n_frames, label):
    cams, status = b.cameras()
    assert not status.any(), status
    for s, stream in enumerate(streams):
        o = run_oracle(kind, quat_mode, stream, n_frames)
        order = [k for k, _ in sorted(o.landmarks.items(), key=lambda kv: kv[1])]
        assert [int(i) for i in b.landmark_ids(s)] == order
        floor = sensitivity_floor(kind, quat_mode, stream.ids[:n_frames], stream.poses[:n_frames]) if s == 0 else None
        assert_parity(f"batch wide {label} seq {s}", b.state(s), b.uncertainty(s), o.state, o.uncertainty, TOL,
                      floor=floor)
        assert rel_err(cams[s], o.state[:10]) < TOL


@pytest.mark.parametrize("kind,quat_mode,n_markers,n_frames,p", [
    ("rot", "scalar_first", 6, 120, 0.6),      # one chunk per frame
    ("ekf", "scalar_first", 10, 200, 0.6),
    ("ekf", "as_written", 10, 40, 0.6),
    ("rot", "scalar_first", 32, 40, 0.8),      # k ~ 26 > 16: two balanced chunks per frame (C2-ROT shape, N = 330)
    ("ekf", "scalar_first", 40, 30, 0.9),      # up to three chunks
])
def test_wide_matches_oracle(kind, quat_mode, n_markers, n_frames, p):
    from aruco_slam_b200 import FilterBatch, synth
    n_seq = 5
    streams = [synth.make_stream(300 + s, n_markers, n_frames, p, kind) for s in range(n_seq)]
    # ragged input: empty frames (no detection => no predict), a repeated id, a sequence that ends early
    streams[1].ids[3] = streams[1].ids[3][:0]
    streams[1].poses[3] = streams[1].poses[3][:0]
    streams[2].ids[5] = np.concatenate([streams[2].ids[5], streams[2].ids[5][:1]])
    streams[2].poses[5] = np.vstack([streams[2].poses[5], streams[2].poses[5][:1] + 1e-3])
    for t in range(n_frames - 4, n_frames):
        streams[3].ids[t] = streams[3].ids[t][:0]
        streams[3].poses[t] = streams[3].poses[t][:0]
    ids, poses, off = pack(streams, n_frames)
    b = FilterBatch(kind, n_seq, n_markers, n_markers + 1, quat_mode=quat_mode, engine="wide")
    assert b.engine == "wide"
    b.upload(ids, poses, off)
    b.run(0, n_frames // 2)
    b.run(n_frames // 2, n_frames)   # state (and a pending update) persists across runs
    b.synchronize()
    _check(b, streams, kind, quat_mode, n_frames, f"{kind}/{quat_mode} {n_markers} markers {n_frames} frames")


def test_wide_is_the_default_for_rotations_at_c2_and_c5_shapes():
    """EKF_Rotations at 32 markers (N = 330) and 64 markers (N = 650, p = 0.25: BASELINE configs[4]) does not fit
    an SM's shared memory: the handle picks the wide engine by itself.  200+ frames at the C5 shape."""
    from aruco_slam_b200 import FilterBatch, synth
    for n_markers, p, n_frames in ((32, 0.5, 60), (64, 0.25, 200)):
        n_seq = 3
        streams = [synth.make_stream(70 + s, n_markers, n_frames, p, "rot") for s in range(n_seq)]
        ids, poses, off = pack(streams, n_frames)
        b = FilterBatch("rot", n_seq, n_markers, n_markers, quat_mode="scalar_first")
        assert b.engine == "wide"
        b.upload(ids, poses, off)
        b.run()
        _check(b, streams, "rot", "scalar_first", n_frames, f"rot {n_markers} markers p={p} {n_frames} frames")


def test_wide_equals_resident_engine():
    """Two independent implementations of the same update behind one handle."""
    from aruco_slam_b200 import FilterBatch, synth
    n_seq, n_markers, n_frames = 8, 12, 80
    ids, poses, off = synth.make_batch(n_seq, n_markers, n_frames, 0.7, "ekf")
    out = {}
    for engine in ("resident", "wide"):
        b = FilterBatch("ekf", n_seq, n_markers, n_markers, quat_mode="scalar_first", engine=engine)
        assert b.engine == engine
        b.upload(ids, poses, off)
        b.run()
        out[engine] = [(b.state(s), b.uncertainty(s)) for s in range(n_seq)]
    for (s1, p1), (s2, p2) in zip(out["resident"], out["wide"]):
        assert rel_err(s1, s2) < 1e-10
        assert rel_err(p1, p2) < 1e-10
    # switching the engine AFTER the upload (C ABI order): the filters restart, the uploaded stream stays
    from aruco_slam_b200 import _lib
    _lib.check(b._lib.asb_batch_set_engine(b._h, 1))
    assert b.engine == "resident"
    b.run()
    for s, (s1, p1) in enumerate(out["resident"]):
        assert np.array_equal(b.state(s), s1) and np.array_equal(b.uncertainty(s), p1)
    _lib.check(b._lib.asb_batch_set_engine(b._h, 2))
    b.run()
    for s, (s2, p2) in enumerate(out["wide"]):
        assert np.array_equal(b.state(s), s2) and np.array_equal(b.uncertainty(s), p2)


def test_wide_capacity_flag_and_reset():
    from aruco_slam_b200 import FilterBatch, synth
    s = synth.make_stream(1, 10, 5, 1.0, "rot")
    ids, poses, off = pack([s], 5)
    b = FilterBatch("rot", 1, 4, 16, engine="wide")   # 10 markers > max_landmarks 4
    b.upload(ids, poses, off)
    b.run()
    _, status = b.cameras()
    assert status[0] == -4
    # reset brings the constructor state back (and clears the flag)
    b.reset()
    cams, status = b.cameras()
    assert status[0] == 0 and np.array_equal(cams[0], [0, 0, 0, 1, 0, 0, 0, 0, 0, 0])


def test_wide_full_size_c2_rot_properties():
    """BASELINE configs[1] with EKF_Rotations at full size (4096 sequences x 32 markers x 1000 frames, N = 330; 64
    distinct streams read through asb_batch_upload_tiled as in bench.py's C5 legs): every sequence finite and
    status 0; the 64 copies of one stream bit-identical (deterministic, sequences do not interact: a race between
    CTAs or stream groups would break it); first and last sequence within 1e-9 of the free-running oracle over all
    1000 frames, with the sensitivity floor beside it."""
    from aruco_slam_b200 import FilterBatch, synth
    n_seq, n_markers, n_frames, unique = 4096, 32, 1000, 64
    ids, poses, off = synth.make_batch(unique, n_markers, n_frames, 0.5, "rot", unique=unique)
    b = FilterBatch("rot", n_seq, n_markers, n_markers)
    assert b.engine == "wide"
    b.upload(ids, poses, off, n_streams=unique)
    b.run()
    cams, status = b.cameras()
    assert not status.any() and np.isfinite(cams).all()
    for s in range(unique):
        assert np.array_equal(cams[s::unique], np.broadcast_to(cams[s], cams[s::unique].shape)), s
    for s in (0, n_seq - 1):
        stream = synth.make_stream(s % unique, n_markers, n_frames, 0.5, "rot")
        o = run_oracle("rot", "scalar_first", stream, n_frames)
        floor = sensitivity_floor("rot", "scalar_first", stream.ids, stream.poses)
        assert_parity(f"batch wide C2-ROT full size (4096 x 32 x 1000) seq {s}", b.state(s), b.uncertainty(s), o.state,
                      o.uncertainty, TOL, floor=floor)
