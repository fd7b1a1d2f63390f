"""CPU: pin oracle/ippe_oracle.py against cv2.solvePnP(IPPE_SQUARE) — golden
vectors made through the reference's own estimate_pose_of_markers
(base_filter.py:92-171), plus a live cross-check when cv2 is importable."""
import numpy as np
import pytest

from conftest import GOLDEN
from oracle import ippe_oracle as io

TOL = 1e-9


def _check(out, ref):
    et = np.abs(out[:, :3] - ref[:, :3]).max(1) / np.abs(ref[:, :3]).max(1)
    er = np.abs(io.rodrigues_matrix(out[:, 3:]) - io.rodrigues_matrix(ref[:, 3:])) \
        .reshape(len(out), -1).max(1)
    return et, er


def test_golden_pose():
    g = np.load(GOLDEN / "ippe_golden.npz")
    out = io.estimate_pose_of_markers(g["corners"], g["camera_matrix"], g["dist_coeffs"],
                                      float(g["marker_size"]))
    et, er = _check(out, g["pose"])
    assert et.max() < TOL and er.max() < TOL  # 0 ordering mismatches on 2000 sets


def test_golden_both_solutions():
    g = np.load(GOLDEN / "ippe_golden.npz")
    sol = io.ippe_square_both(g["corners"], g["camera_matrix"], g["dist_coeffs"],
                              float(g["marker_size"]))
    first = np.hstack([sol["t1"], io.rodrigues_vec(sol["R1"])])
    second = np.hstack([sol["t2"], io.rodrigues_vec(sol["R2"])])
    for out, ref in ((first, g["both"][:, 0]), (second, g["both"][:, 1])):
        et, er = _check(out, ref)
        # the second solution of a near-frontal marker is ill-conditioned; 1e-7 there
        assert np.median(et) < 1e-12 and et.max() < 1e-7 and er.max() < 1e-7


def test_live_cv2(calibration):
    cv2 = pytest.importorskip("cv2")
    from aruco_slam_b200 import synth
    km, dist = calibration
    corners = synth.make_corner_sets(500, km, dist, seed=5)
    ms = 0.16
    mp = np.array([[-ms / 2, ms / 2, 0], [ms / 2, ms / 2, 0], [ms / 2, -ms / 2, 0],
                   [-ms / 2, -ms / 2, 0]], dtype=np.float32)
    ref = []
    for c in corners:
        _, r, t = cv2.solvePnP(mp, c.reshape(1, 4, 2), km, dist, flags=cv2.SOLVEPNP_IPPE_SQUARE)
        ref.append(np.hstack([t.ravel(), r.ravel()]))
    out = io.estimate_pose_of_markers(corners, km, dist, ms)
    et, er = _check(out, np.array(ref))
    assert et.max() < TOL and er.max() < TOL
    und = cv2.undistortPoints(corners.reshape(-1, 1, 2), km, dist).reshape(-1, 4, 2)
    x, y = io.undistort_normalize(corners, km, dist)
    assert np.array_equal(und[..., 0].astype(np.float64), x)  # float32 rounding replicated
    assert np.array_equal(und[..., 1].astype(np.float64), y)
