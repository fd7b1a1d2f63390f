"""Many independent filters advanced together (BASELINE configs C2 / C5).  Each sequence behaves
exactly like one ``EKF`` / ``EKF_Rotations`` instance fed its own frames through ``observe``
(extended_kalman_filter.py:58-82, ekf_with_rotations.py:66-90).  Two engines behind one handle:
``resident`` (one CTA per sequence, the filter in shared memory for all frames) and ``wide``
(frame-synchronous kernels over all sequences, covariance tiled in HBM: EKF_Rotations at 32 / 64
markers); ``auto`` picks resident when the covariance fits an SM."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

CAM_DIMS = 10


class FilterBatch:
    def __init__(self, kind: str, n_seq: int, max_landmarks: int, max_obs_per_frame: int, *,
                 quat_mode: str = "as_written", initial_camera_pose=None, noise: dict | None = None,
                 device: int = 0, engine: str = "auto") -> None:
        kinds = {"ekf": _lib.KIND_EKF, "rot": _lib.KIND_ROT, "ekf_rotations": _lib.KIND_ROT}
        modes = {"as_written": _lib.QUAT_AS_WRITTEN, "scalar_first": _lib.QUAT_SCALAR_FIRST}
        self.kind = kinds[kind]
        self.lm_dims = 3 if self.kind == _lib.KIND_EKF else 10
        self.n_seq = int(n_seq)
        self._lib = _lib.lib()
        self._h = None
        pose = _lib.as_f64([0, 0, 0, 1, 0, 0, 0, 0, 0, 0] if initial_camera_pose is None
                           else initial_camera_pose, (CAM_DIMS,))
        cfg = None
        if noise is not None:
            nz = _lib.Noise()
            _lib.check(self._lib.asb_default_noise(self.kind, C.byref(nz)))
            for key, val in noise.items():
                setattr(nz, key, float(val))
            cfg = C.byref(nz)
        h = C.c_void_p()
        _lib.check(self._lib.asb_batch_create(self.kind, modes[quat_mode], _lib.dptr(pose), cfg,
                                              self.n_seq, int(max_landmarks), int(max_obs_per_frame),
                                              int(device), C.byref(h)))
        self._h = h
        self.n_frames = 0
        self._keep = None
        if engine != "auto":
            _lib.check(self._lib.asb_batch_set_engine(self._h, {"resident": 1, "wide": 2}[engine]))

    @property
    def engine(self) -> str:
        e = C.c_int()
        _lib.check(self._lib.asb_batch_get_engine(self._h, C.byref(e)))
        return {1: "resident", 2: "wide"}[e.value]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.asb_batch_destroy(h)
            except Exception:  # interpreter shutdown
                pass

    def reset(self) -> None:
        """Back to freshly constructed filters (asynchronous)."""
        _lib.check(self._lib.asb_batch_reset(self._h))

    def set_stream(self, cuda_stream: int | None) -> None:
        _lib.check(self._lib.asb_batch_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def upload(self, ids, poses, offsets, n_streams: int | None = None) -> None:
        """Copy the packed observations of all sequences to the device (async on the
        stream; the arrays are kept alive until the next upload).  ``n_streams``: the arrays hold
        only that many distinct streams and sequence ``s`` reads stream ``s % n_streams``."""
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        off = np.ascontiguousarray(np.asarray(offsets, dtype=np.int64))
        if ids.size != poses.shape[0] or off[-1] != ids.size:
            raise ValueError("ids, poses and offsets disagree on the number of observations")
        ns = self.n_seq if n_streams is None else int(n_streams)
        n_frames = (off.size - 1) // ns
        if n_frames * ns + 1 != off.size:
            raise ValueError("offsets must have n_streams * n_frames + 1 entries")
        self._keep = (ids, poses, off)
        if n_streams is None:
            _lib.check(self._lib.asb_batch_upload(self._h, _lib.iptr(ids), _lib.dptr(poses),
                                                  _lib.lptr(off), n_frames))
        else:
            _lib.check(self._lib.asb_batch_upload_tiled(self._h, _lib.iptr(ids), _lib.dptr(poses),
                                                        _lib.lptr(off), n_frames, ns))
        self.n_frames = n_frames

    def run(self, t0: int = 0, t1: int | None = None) -> None:
        """``observe`` for frames [t0, t1) of every sequence (asynchronous)."""
        _lib.check(self._lib.asb_batch_run(self._h, int(t0),
                                           int(self.n_frames if t1 is None else t1)))

    def synchronize(self) -> None:
        _lib.check(self._lib.asb_batch_synchronize(self._h))

    def dims(self, seq: int):
        nl, ns = C.c_int(), C.c_int()
        _lib.check(self._lib.asb_batch_get_dims(self._h, int(seq), C.byref(nl), C.byref(ns)))
        return nl.value, ns.value

    def state(self, seq: int) -> np.ndarray:
        _, n = self.dims(seq)
        out = np.empty(n)
        _lib.check(self._lib.asb_batch_get_state(self._h, int(seq), _lib.dptr(out), n))
        return out

    def uncertainty(self, seq: int) -> np.ndarray:
        _, n = self.dims(seq)
        out = np.empty((n, n))
        _lib.check(self._lib.asb_batch_get_cov(self._h, int(seq), _lib.dptr(out), n))
        return out

    def load_checkpoint(self, seq: int, state, uncertainty, ids_in_index_order) -> None:
        """Replace one sequence's filter (id map, state, covariance): checkpoint restore / per-frame
        re-synchronisation in the parity tests, the batch counterpart of ``_CudaFilter.load_checkpoint``."""
        ids = _lib.as_i32(ids_in_index_order)
        st = _lib.as_f64(state).reshape(-1)
        cov = _lib.as_f64(uncertainty)
        _lib.check(self._lib.asb_batch_set_landmarks(self._h, int(seq), _lib.iptr(ids), ids.size))
        _lib.check(self._lib.asb_batch_set_state(self._h, int(seq), _lib.dptr(st), st.size))
        _lib.check(self._lib.asb_batch_set_cov(self._h, int(seq), _lib.dptr(cov), cov.shape[0]))

    def landmark_ids(self, seq: int) -> np.ndarray:
        nl, _ = self.dims(seq)
        out = np.empty(max(nl, 1), dtype=np.int32)
        _lib.check(self._lib.asb_batch_get_landmark_ids(self._h, int(seq), _lib.iptr(out), out.size))
        return out[:nl]

    def cameras(self):
        """((n_seq,10) camera poses, (n_seq,) status codes)."""
        out = np.empty((self.n_seq, CAM_DIMS))
        st = np.empty(self.n_seq, dtype=np.int32)
        _lib.check(self._lib.asb_batch_get_cameras(self._h, _lib.dptr(out), _lib.iptr(st)))
        return out, st

    def get_poses(self, seq: int):
        s = self.state(seq)
        return s[:CAM_DIMS], s[CAM_DIMS:].reshape(-1, self.lm_dims)
