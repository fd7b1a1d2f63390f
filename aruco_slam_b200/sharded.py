"""One EKF / EKF_Rotations whose covariance rows are spread over the GPUs of a box, one
process per GPU (BASELINE config C3 at N GPUs).

Same algorithm and same numbers as ``EKF`` / ``EKF_Rotations`` (``observe`` =
extended_kalman_filter.py:58-82 / ekf_with_rotations.py:66-90).  The state and the id -> index
map are replicated; rank g owns rows [row0, row1) of ``uncertainty``.  Every rank calls the
same methods with the same arguments; each update runs two NCCL collectives inside the
library (filter_shard.cu).  ``torch.distributed`` is only the plumbing that carries the NCCL
unique id to the ranks and, for the ``uncertainty`` getter, gathers the row blocks.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch  # noqa: F401 — loads torch's bundled libnccl.so.2 BEFORE the library opens NCCL (one copy per process)

from . import _lib

CAM_DIMS = 10
_KINDS = {"ekf": _lib.KIND_EKF, "rot": _lib.KIND_ROT, "ekf_rotations": _lib.KIND_ROT}
_MODES = {"as_written": _lib.QUAT_AS_WRITTEN, "scalar_first": _lib.QUAT_SCALAR_FIRST}


def partition(kind: str, max_landmarks: int, world: int, rank: int) -> tuple[int, int, int]:
    """(row0, row1, ld) of `rank`: whole 64-column panel blocks, the same number on every rank."""
    r0, r1, ld = C.c_int(), C.c_int(), C.c_int()
    _lib.check(_lib.lib().asb_shard_partition(_KINDS[kind], int(max_landmarks), int(world), int(rank),
                                              C.byref(r0), C.byref(r1), C.byref(ld)))
    return r0.value, r1.value, ld.value


def exchange_unique_id(dist, rank: int) -> bytes:
    """ncclGetUniqueId on rank 0, broadcast through torch.distributed (any backend)."""
    box = [None]
    if rank == 0:
        buf = C.create_string_buffer(128)
        _lib.check(_lib.lib().asb_shard_unique_id(buf, 128))
        box[0] = bytes(buf.raw)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast_object_list(box, src=0)
    return box[0]


class ShardedFilter:
    def __init__(self, kind: str, initial_camera_pose, max_landmarks: int, *, dist=None, rank: int = 0,
                 world: int = 1, device: int = 0, quat_mode: str = "as_written", noise: dict | None = None,
                 unique_id: bytes | None = None) -> None:
        self.kind = _KINDS[kind]
        self.lm_dims = 3 if self.kind == _lib.KIND_EKF else 10
        self.rank, self.world, self._dist = int(rank), int(world), dist
        self._lib = _lib.lib()
        self._h = None
        self.row0, self.row1, self.ld = partition(kind, max_landmarks, world, rank)
        uid = unique_id if unique_id is not None else exchange_unique_id(dist, rank)
        pose = _lib.as_f64(initial_camera_pose, (CAM_DIMS,))
        cfg = None
        if noise is not None:
            nz = _lib.Noise()
            _lib.check(self._lib.asb_default_noise(self.kind, C.byref(nz)))
            for key, val in noise.items():
                setattr(nz, key, float(val))
            cfg = C.byref(nz)
        h = C.c_void_p()
        self._uid = C.create_string_buffer(uid, 128)
        _lib.check(self._lib.asb_shard_create(self.kind, _MODES[quat_mode], _lib.dptr(pose), cfg,
                                              int(max_landmarks), int(device), self.rank, self.world,
                                              self._uid, C.byref(h)))
        self._h = h
        self._keep = None

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.asb_shard_destroy(h)
            except Exception:  # interpreter shutdown
                pass

    def close(self) -> None:
        h, self._h = self._h, None
        if h:
            _lib.check(self._lib.asb_shard_destroy(h))

    def set_stream(self, cuda_stream: int | None) -> None:
        _lib.check(self._lib.asb_shard_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def synchronize(self) -> None:
        _lib.check(self._lib.asb_shard_synchronize(self._h))

    def observe(self, ids, poses) -> None:
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        if ids.size != poses.shape[0]:
            raise ValueError(f"{ids.size} ids but {poses.shape[0]} poses")
        _lib.check(self._lib.asb_shard_observe(self._h, _lib.iptr(ids), _lib.dptr(poses), int(ids.size)))

    def stream_upload(self, ids, poses, offsets) -> None:
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        off = np.ascontiguousarray(np.asarray(offsets, dtype=np.int64))
        _lib.check(self._lib.asb_shard_stream_upload(self._h, _lib.iptr(ids), _lib.dptr(poses), _lib.lptr(off),
                                                     int(off.size - 1)))

    def stream_run(self, t0: int, t1: int) -> None:
        _lib.check(self._lib.asb_shard_stream_run(self._h, int(t0), int(t1)))

    def _dims(self):
        nl, ns = C.c_int(), C.c_int()
        _lib.check(self._lib.asb_shard_get_dims(self._h, C.byref(nl), C.byref(ns)))
        return nl.value, ns.value

    @property
    def num_landmarks(self) -> int:
        return self._dims()[0]

    @property
    def state(self) -> np.ndarray:
        _, n = self._dims()
        out = np.empty(n)
        _lib.check(self._lib.asb_shard_get_state(self._h, _lib.dptr(out), n))
        return out

    def get_poses(self):
        s = self.state
        return s[:CAM_DIMS], s[CAM_DIMS:].reshape(-1, self.lm_dims)

    def local_rows(self) -> np.ndarray:
        """This rank's live rows of ``uncertainty``: (rows, N)."""
        _, n = self._dims()
        rows = max(0, min(self.row1, n) - self.row0)
        out = np.empty((max(rows, 1), n))
        got = C.c_int()
        _lib.check(self._lib.asb_shard_get_cov_rows(self._h, _lib.dptr(out), out.size, C.byref(got)))
        return out[:got.value]

    @property
    def uncertainty(self) -> np.ndarray:
        """The full (N, N) covariance on every rank (row blocks gathered through torch.distributed)."""
        mine = self.local_rows()
        if self.world == 1:
            return mine
        parts = [None] * self.world
        self._dist.all_gather_object(parts, mine)
        return np.concatenate([p for p in parts if p.shape[0]], axis=0)

    @property
    def landmarks(self) -> dict:
        nl, _ = self._dims()
        out = np.empty(max(nl, 1), dtype=np.int32)
        _lib.check(self._lib.asb_shard_get_landmark_ids(self._h, _lib.iptr(out), out.size))
        return {int(i): k for k, i in enumerate(out[:nl])}
