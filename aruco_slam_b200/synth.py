"""Deterministic synthetic observation streams.

The reference ships no input video, so every parity test and benchmark is
driven by these streams: a camera on a smooth closed path looking at static
markers, with the marker poses delivered exactly as
``BaseFilter.estimate_pose_of_markers`` would deliver them
(/root/reference/filters/base_filter.py:168-171): one ``(k, 6)`` float64 array
``[t(3), rvec(3)]`` per frame plus the ``(k,)`` integer ids.

numpy only (PCG64 ``default_rng``); sequence ``s`` uses ``seed = 1234 + s``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


def euler_xyz_matrix(a: np.ndarray) -> np.ndarray:
    """scipy ``Rotation.from_euler('xyz', a).as_matrix()`` (extrinsic x-y-z):
    R = Rz(a2) @ Ry(a1) @ Rx(a0)."""
    cx, sx = np.cos(a[0]), np.sin(a[0])
    cy, sy = np.cos(a[1]), np.sin(a[1])
    cz, sz = np.cos(a[2]), np.sin(a[2])
    rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return rz @ ry @ rx


def matrix_to_euler_xyz(r: np.ndarray) -> np.ndarray:
    """Inverse of :func:`euler_xyz_matrix` away from gimbal lock."""
    a1 = -np.arcsin(np.clip(r[2, 0], -1.0, 1.0))
    a0 = np.arctan2(r[2, 1], r[2, 2])
    a2 = np.arctan2(r[1, 0], r[0, 0])
    return np.array([a0, a1, a2])


@dataclass
class Stream:
    """One synthetic sequence.  ``ids``/``poses`` are per-frame lists."""
    ids: list            # T arrays (k_t,) int32
    poses: list          # T arrays (k_t, 6) float64
    landmarks: np.ndarray  # (n, 3) ground-truth map positions
    n_markers: int

    @property
    def n_frames(self) -> int:
        return len(self.ids)

    @property
    def n_updates(self) -> int:
        return int(sum(len(i) for i in self.ids))

    def packed(self):
        """CSR-style packing: (ids (M,), poses (M,6), frame_offsets (T+1,))."""
        off = np.zeros(self.n_frames + 1, dtype=np.int32)
        off[1:] = np.cumsum([len(i) for i in self.ids])
        ids = np.concatenate(self.ids).astype(np.int32) if self.n_updates else np.zeros(0, np.int32)
        poses = np.concatenate(self.poses).astype(np.float64) if self.n_updates \
            else np.zeros((0, 6))
        return ids, np.ascontiguousarray(poses), off


def camera_pose(frame: int, period: int = 1000):
    tau = 2.0 * np.pi * frame / period
    c = np.array([0.5 * np.sin(tau), 0.2 * np.sin(2 * tau), 0.5 * (1 - np.cos(tau))])
    r_mc = euler_xyz_matrix(np.array([0.1 * np.sin(tau), 0.25 * np.sin(tau),
                                      0.05 * np.sin(3 * tau)]))
    return c, r_mc


def make_stream(seq: int = 0, n_markers: int = 10, n_frames: int = 1000, p_visible: float = 0.8,
                kind: str = "ekf", wide_map: bool = False, period: int = 1000,
                id_base: int = 0) -> Stream:
    """Synthetic stream per SURVEY.md §8(d).

    kind='ekf': rvec slot is N(0, 0.1^2) noise (EKF ignores it,
    extended_kalman_filter.py:192).  kind='rot': rvec slot carries
    euler_xyz(R_mc^T R_ml) + N(0, 0.01^2), which EKF_Rotations reads as Euler
    angles (ekf_with_rotations.py:216-219, 307-310).
    All markers are visible in frame 0 (rotation is identity there).
    """
    rng = np.random.default_rng(1234 + seq)
    if wide_map:
        lo = np.array([-20.0, -5.0, 2.0])
        hi = np.array([20.0, 5.0, 40.0])
        lm = rng.uniform(lo, hi, size=(n_markers, 3))
    else:
        lm = rng.uniform(-2.0, 2.0, size=(n_markers, 3)) + np.array([0.0, 0.0, 5.0])
    r_ml = [euler_xyz_matrix(rng.normal(0.0, 0.2, size=3)) for _ in range(n_markers)]
    r_ml_arr = np.stack(r_ml)
    marker_ids = id_base + rng.permutation(n_markers).astype(np.int32)  # id != index on purpose

    ids_out, poses_out = [], []
    for f in range(n_frames):
        c, r_mc = camera_pose(f, period)
        if f == 0:
            vis = np.ones(n_markers, dtype=bool)
        else:
            vis = rng.random(n_markers) < p_visible
            if not vis.any():
                vis[rng.integers(n_markers)] = True
        idx = np.flatnonzero(vis)
        rng.shuffle(idx)
        k = len(idx)
        t = (lm[idx] - c) @ r_mc + rng.normal(0.0, 0.01, size=(k, 3))  # R_mc^T (l - c)
        if kind == "rot":
            rr = np.matmul(r_mc.T, r_ml_arr[idx])            # (k,3,3) R_mc^T R_ml, then euler_xyz of each
            rv = np.stack([np.arctan2(rr[:, 2, 1], rr[:, 2, 2]), -np.arcsin(np.clip(rr[:, 2, 0], -1.0, 1.0)),
                           np.arctan2(rr[:, 1, 0], rr[:, 0, 0])], axis=1) + rng.normal(0.0, 0.01, size=(k, 3))
        else:
            rv = rng.normal(0.0, 0.1, size=(k, 3))
        ids_out.append(marker_ids[idx].astype(np.int32))
        poses_out.append(np.hstack([t, rv]))
    return Stream(ids_out, poses_out, lm, n_markers)


def _packed_stream(seq, n_markers, n_frames, p_visible, kind):
    return make_stream(seq, n_markers, n_frames, p_visible, kind).packed()


def make_batch(n_seq: int, n_markers: int, n_frames: int, p_visible: float, kind: str = "ekf",
               seq_base: int = 0, unique: int | None = None, workers: int = 0):
    """Pack ``n_seq`` sequences for the batched engine.

    Returns (ids (M,) i32, poses (M,6) f64, frame_offsets (n_seq*(T)+1,) i32 with
    the frames of sequence s occupying rows [s*T, (s+1)*T]).  Generating 4096
    distinct 1000-frame sequences in numpy takes minutes, so ``unique`` distinct
    sequences are generated and tiled (the bench says so in ``data``); parity
    tests always use ``unique=None`` (all distinct).
    """
    u = n_seq if unique is None else min(unique, n_seq)
    if workers and workers > 1 and u >= 4 * workers:
        import multiprocessing as mp
        with mp.get_context("spawn").Pool(workers) as pool:
            packs = pool.starmap(_packed_stream, [(seq_base + s, n_markers, n_frames, p_visible, kind)
                                                  for s in range(u)], chunksize=max(1, u // (8 * workers)))
    else:
        packs = [_packed_stream(seq_base + s, n_markers, n_frames, p_visible, kind) for s in range(u)]
    ids_l, poses_l, counts = [], [], []
    for s in range(n_seq):
        i, p, off = packs[s % u]
        ids_l.append(i)
        poses_l.append(p)
        counts.append(np.diff(off))
    ids = np.concatenate(ids_l).astype(np.int32)
    poses = np.ascontiguousarray(np.concatenate(poses_l))
    offsets = np.zeros(n_seq * n_frames + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(np.concatenate(counts))
    return ids, poses, offsets


def make_map_build_stream(n_markers: int = 2000, per_frame: int = 20, timed_frames: int = 200,
                          seed: int = 1234):
    """C3 stream: build frames each observe ``per_frame`` NEW
    markers (ids 20f..20f+19), then ``timed_frames`` frames each observe exactly
    ``per_frame`` distinct known markers drawn uniformly."""
    rng = np.random.default_rng(seed)
    lo = np.array([-20.0, -5.0, 2.0])
    hi = np.array([20.0, 5.0, 40.0])
    lm = rng.uniform(lo, hi, size=(n_markers, 3))
    n_build = n_markers // per_frame
    ids_out, poses_out = [], []
    for f in range(n_build + timed_frames):
        # the camera stays near the origin with small rotation during the build
        c, r_mc = camera_pose(f % 1000, 1000)
        if f < n_build:
            idx = np.arange(f * per_frame, (f + 1) * per_frame)
        else:
            idx = rng.choice(n_markers, size=per_frame, replace=False)
        t = (lm[idx] - c) @ r_mc + rng.normal(0.0, 0.01, size=(len(idx), 3))
        rv = rng.normal(0.0, 0.1, size=(len(idx), 3))
        ids_out.append(idx.astype(np.int32))
        poses_out.append(np.hstack([t, rv]))
    return Stream(ids_out, poses_out, lm, n_markers), n_build


def reference_calibration():
    """The reference's calibration/*.npy (data, SURVEY.md §2 "IN (data)"), shipped
    with the package so the GPU box (no /root/reference) uses the same intrinsics."""
    from pathlib import Path
    d = Path(__file__).resolve().parent / "calibration"
    return np.load(d / "camera_matrix.npy"), np.load(d / "dist_coeffs.npy")


def project_points(obj, rvec, tvec, camera_matrix, dist_coeffs):
    """Pin-hole + 5-coefficient Brown-Conrady projection (the forward model of
    ``cv2.projectPoints``) used to synthesise corner sets.  obj (4,3); rvec/tvec (B,3)."""
    rvec = np.atleast_2d(rvec)
    tvec = np.atleast_2d(tvec)
    th = np.sqrt((rvec * rvec).sum(1))
    k = rvec / np.where(th > 0, th, 1.0)[:, None]
    kx = np.zeros((rvec.shape[0], 3, 3))
    kx[:, 0, 1], kx[:, 0, 2] = -k[:, 2], k[:, 1]
    kx[:, 1, 0], kx[:, 1, 2] = k[:, 2], -k[:, 0]
    kx[:, 2, 0], kx[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    rm = np.eye(3)[None] + s * kx + (1 - c) * (kx @ kx)
    pc = np.einsum("bij,pj->bpi", rm, obj) + tvec[:, None, :]
    x = pc[..., 0] / pc[..., 2]
    y = pc[..., 1] / pc[..., 2]
    k1, k2, p1, p2, k3 = np.asarray(dist_coeffs, dtype=np.float64).ravel()[:5]
    r2 = x * x + y * y
    cd = 1 + ((k3 * r2 + k2) * r2 + k1) * r2
    xd = x * cd + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
    yd = y * cd + p1 * (r2 + 2 * y * y) + 2 * p2 * x * y
    km = np.asarray(camera_matrix, dtype=np.float64)
    u = km[0, 0] * xd + km[0, 2]
    v = km[1, 1] * yd + km[1, 2]
    return np.stack([u, v], axis=-1), pc[..., 2]


def make_corner_sets(n_sets: int, camera_matrix, dist_coeffs, marker_size: float = 0.16,
                     seed: int = 4321, pixel_noise: float = 0.3):
    """C4: random marker poses in front of the camera, projected,
    + N(0, 0.3^2) px noise, cast to float32, kept if inside 1920x1080."""
    rng = np.random.default_rng(seed)
    h = marker_size / 2
    obj = np.array([[-h, h, 0], [h, h, 0], [h, -h, 0], [-h, -h, 0]], dtype=np.float64)
    out = []
    have = 0
    while have < n_sets:
        b = max(1024, int((n_sets - have) * 1.6))
        rvec = rng.normal(0.0, 0.5, size=(b, 3))
        tvec = np.stack([rng.uniform(-1.5, 1.5, b), rng.uniform(-0.8, 0.8, b),
                         rng.uniform(0.8, 8.0, b)], axis=1)
        uv, z = project_points(obj, rvec, tvec, camera_matrix, dist_coeffs)
        uv = uv + rng.normal(0.0, pixel_noise, size=uv.shape)
        ok = (z > 0.1).all(1) & (uv[..., 0] > 0).all(1) & (uv[..., 0] < 1920).all(1) \
            & (uv[..., 1] > 0).all(1) & (uv[..., 1] < 1080).all(1)
        out.append(uv[ok].astype(np.float32))
        have += int(ok.sum())
    return np.ascontiguousarray(np.concatenate(out)[:n_sets])
