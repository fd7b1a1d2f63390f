"""Generate tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE in the build
container (python oracle/make_golden.py).  TEST INFRASTRUCTURE.

The reference has no tests or golden vectors, so these fixtures —
outputs of the reference's own classes (imported by oracle/ref_import.py) and of
``cv2.solvePnP`` called exactly as base_filter.py:156-164 calls it — are what
pins oracle/ekf_oracle.py and oracle/ippe_oracle.py.  /root/reference does not
exist on the GPU box; only the committed .npz files travel.

Fixtures:
  ekf_steps.npz   one-step goldens (pre state/P/landmarks, ids, poses ->
                  z, h, H, post state/P) for EKF as-written, EKF scalar-first and
                  EKF_Rotations, taken along the reference's own free run.
  stream_<mode>.npz  free-running C1 trajectory (10 markers, p=0.8): inputs,
                  per-frame camera pose, final state / P / landmark order.
  ippe_golden.npz corner sets -> cv2.solvePnP pose, both solvePnPGeneric solutions.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from aruco_slam_b200 import synth  # noqa: E402
from oracle import ref_import  # noqa: E402

GOLD = ROOT / "tests" / "golden"


def late_marker_stream(kind: str, n_frames: int):
    """12 markers; markers 8..11 are hidden before frame 30 (mid-stream
    augmentation while the camera is rotated), frame 40 repeats an
    id twice (duplicate row blocks)."""
    s = synth.make_stream(7, 12, n_frames, 0.7, kind)
    late = set(np.sort(s.ids[0])[8:].tolist())
    for f in range(min(30, n_frames)):
        keep = np.array([i not in late for i in s.ids[f]])
        if not keep.any():
            keep[0] = True
        s.ids[f] = s.ids[f][keep]
        s.poses[f] = s.poses[f][keep]
    if n_frames > 40:
        s.ids[40] = np.concatenate([s.ids[40], s.ids[40][:1]])
        s.poses[40] = np.vstack([s.poses[40], s.poses[40][:1] + 1e-3])
    return s


def make_filter(mode: str):
    if mode == "ekf_as_written":
        return ref_import.make_ekf(quat_mode="as_written")
    if mode == "ekf_scalar_first":
        return ref_import.make_ekf(quat_mode="scalar_first")
    if mode == "rot":
        return ref_import.make_rot()
    raise ValueError(mode)


def lm_order(f):
    """ids in index order."""
    inv = {v: k for k, v in f.landmarks.items()}
    return np.array([int(inv[i]) for i in range(f.num_landmarks)], dtype=np.int64)


def step_goldens():
    out = {}
    for mode in ("ekf_as_written", "ekf_scalar_first", "rot"):
        kind = "rot" if mode == "rot" else "ekf"
        f = make_filter(mode)
        n_frames = 120
        s = late_marker_stream(kind, n_frames)
        want = {0, 1, 2, 29, 30, 31, 40, 41, 80, 119}
        for t in range(n_frames):
            ids, poses = s.ids[t], s.poses[t]
            if t in want:
                key = f"{mode}/f{t}"
                out[f"{key}/pre_state"] = np.array(f.state, dtype=np.float64)
                out[f"{key}/pre_P"] = np.array(f.uncertainty, dtype=np.float64)
                out[f"{key}/pre_lm"] = lm_order(f)
                out[f"{key}/ids"] = np.asarray(ids, dtype=np.int64)
                out[f"{key}/poses"] = np.asarray(poses, dtype=np.float64)
                # replay observe() by hand to capture the intermediates
                for idx, pose in zip(ids, poses):
                    if idx in f.landmarks:
                        continue
                    f.add_marker(idx, pose)
                f.predict()
                z, h, dh = f.parse_poses(ids, poses)
                out[f"{key}/z"] = np.asarray(z, dtype=np.float64)
                out[f"{key}/h"] = np.asarray(h, dtype=np.float64)
                out[f"{key}/H"] = np.asarray(dh, dtype=np.float64)
                f.update(ids, poses)
                out[f"{key}/post_state"] = np.array(f.state, dtype=np.float64)
                out[f"{key}/post_P"] = np.array(f.uncertainty, dtype=np.float64)
                out[f"{key}/post_lm"] = lm_order(f)
            else:
                f.observe(ids, poses)
    np.savez_compressed(GOLD / "ekf_steps.npz", **out)
    print("ekf_steps.npz", len(out), "arrays")


def stream_goldens():
    for mode, n_frames in (("ekf_as_written", 60), ("ekf_scalar_first", 1000), ("rot", 1000)):
        kind = "rot" if mode == "rot" else "ekf"
        f = make_filter(mode)
        s = synth.make_stream(0, 10, n_frames, 0.8, kind)
        ids, poses, off = s.packed()
        cam = np.empty((n_frames, 10))
        for t in range(n_frames):
            f.observe(s.ids[t], s.poses[t])
            cam[t] = f.state[:10]
        np.savez_compressed(
            GOLD / f"stream_{mode}.npz", ids=ids, poses=poses, offsets=off, cam=cam,
            final_state=np.array(f.state, dtype=np.float64),
            final_P=np.array(f.uncertainty, dtype=np.float64), lm=lm_order(f))
        print("stream", mode, n_frames, "frames")


def ippe_goldens():
    import cv2
    km, dist = synth.reference_calibration()
    corners = synth.make_corner_sets(2000, km, dist, 0.16, seed=99)
    bf = ref_import.base_filter_for_pnp()
    ids = np.arange(len(corners))
    ref = bf.estimate_pose_of_markers([c.reshape(1, 4, 2) for c in corners], ids, 0.16)
    ms = 0.16
    mp = np.array([[-ms / 2, ms / 2, 0], [ms / 2, ms / 2, 0], [ms / 2, -ms / 2, 0],
                   [-ms / 2, -ms / 2, 0]], dtype=np.float32)
    both = np.empty((len(corners), 2, 6))
    errs = np.empty((len(corners), 2))
    for i, c in enumerate(corners):
        n, rv, tv, er = cv2.solvePnPGeneric(mp, c.reshape(1, 4, 2), km, dist,
                                            flags=cv2.SOLVEPNP_IPPE_SQUARE)
        for j in range(2):
            both[i, j, :3] = tv[j].ravel()
            both[i, j, 3:] = rv[j].ravel()
            errs[i, j] = float(np.ravel(er)[j])
    np.savez_compressed(GOLD / "ippe_golden.npz", corners=corners, camera_matrix=km,
                        dist_coeffs=dist, marker_size=np.float64(0.16), pose=ref,
                        both=both, pixel_err=errs, cv2_version=np.array(cv2.__version__))
    print("ippe_golden.npz", corners.shape)


if __name__ == "__main__":
    GOLD.mkdir(parents=True, exist_ok=True)
    which = sys.argv[1:] or ["steps", "streams", "ippe"]
    if "ippe" in which:
        ippe_goldens()
    if "steps" in which:
        step_goldens()
    if "streams" in which:
        stream_goldens()
