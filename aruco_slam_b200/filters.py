"""Drop-in filter classes with the reference's ``BaseFilter`` API, backed by the CUDA
library through the C ABI (ctypes).

Mirrors (paths relative to the reference tree):
  ``BaseFilter``      filters/base_filter.py:35-381
  ``EKF``             filters/extended_kalman_filter.py:37-357
  ``EKF_Rotations``   filters/ekf_with_rotations.py:43-431
Same constructor positional arguments, method names, argument meaning, return shapes
and error behaviour (``FileNotFoundError`` for missing calibration, ``KeyError`` for an
unknown id in ``update``), so ``main.run_slam``-style loops work unchanged.  State and
covariance live in HBM; ``state`` / ``uncertainty`` are properties that copy on access.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import _lib
from .ippe import estimate_pose_of_markers as _ippe_poses

# base_filter.py:12-13
CALIB_MTX_FILE = "./calibration/camera_matrix.npy"
DIST_COEFFS_FILE = "./calibration/dist_coeffs.npy"

CAM_DIMS = 10
XYZ_DIMS = slice(0, 3)
QUAT_DIMS = slice(3, 7)
ERROR_DIMS = slice(7, 10)

NOT_IMPLEMENTED_ERROR = "This method is not implemented in the base class and should be " \
                        "implemented in a subclass."


class BaseFilter:
    """Front end shared by the filters (base_filter.py:35): intrinsics, marker pose
    estimation (batched IPPE kernel instead of the per-marker solvePnP loop),
    ``process_frame`` glue and ``save_map``."""

    def __init__(self, initial_pose, map_file=None, aruco_dict=None, *, camera_matrix=None,
                 dist_coeffs=None, device: int = 0) -> None:
        if camera_matrix is None or dist_coeffs is None:
            # reference behaviour: CWD-relative files, FileNotFoundError if absent
            # (base_filter.py:55-63)
            if not Path(CALIB_MTX_FILE).exists():
                raise FileNotFoundError("Camera matrix not found. Run calibration.py first.")
            if not Path(DIST_COEFFS_FILE).exists():
                raise FileNotFoundError(
                    "Distortion coefficients not found. Run calibration.py first.")
            camera_matrix = np.load(CALIB_MTX_FILE)
            dist_coeffs = np.load(DIST_COEFFS_FILE)
        self.calib_matrix = np.asarray(camera_matrix, dtype=np.float64)
        self.dist_coeffs = np.asarray(dist_coeffs, dtype=np.float64)
        self._aruco_dict = aruco_dict
        self._detector = None
        self.camera_pose = initial_pose
        self.device = device
        if map_file is not None:
            self.load_map(map_file)

    # -- detection is upstream of the hot path (out of scope); built lazily from cv2 ----
    @property
    def detector(self):
        if self._detector is None:
            self._detector = self.init_aruco_detector(self._aruco_dict)
        return self._detector

    def init_aruco_detector(self, aruco_dict):
        """base_filter.py:74-90 (OpenCV image-space detection; not accelerated here)."""
        import cv2
        if aruco_dict is None:
            aruco_dict = cv2.aruco.DICT_5X5_50
        aruco_dict = cv2.aruco.getPredefinedDictionary(aruco_dict)
        params = cv2.aruco.DetectorParameters()
        params.cornerRefinementMethod = cv2.aruco.CORNER_REFINE_SUBPIX
        params.cornerRefinementWinSize = 3
        params.cornerRefinementMaxIterations = 3
        params.adaptiveThreshWinSizeMin = 3
        params.adaptiveThreshWinSizeMax = 30
        return cv2.aruco.ArucoDetector(aruco_dict, params)

    def estimate_pose_of_markers(self, corners, ids, marker_size: float) -> np.ndarray:
        """(k,6) ``[tvec, rvec]`` float64 rows (base_filter.py:92-171), one batched
        IPPE kernel launch for all k markers."""
        return _ippe_poses(corners, self.calib_matrix, self.dist_coeffs, marker_size,
                           device=self.device)

    def process_detections(self, corners, ids, should_filter=True, iteration=0,
                           marker_size: float = 0.16):
        """``process_frame`` after ``detectMarkers`` (base_filter.py:196-212): the
        headless entry for callers that already hold corners/ids."""
        detected_poses = np.array([])
        if ids is not None and len(ids) > 0:
            ids = np.asarray(ids).flatten()
            if should_filter and hasattr(self, "_observe_corners"):
                # IPPE and the filter update in one call: the poses never leave the device in between
                detected_poses = self._observe_corners(ids, corners, marker_size)
            else:
                detected_poses = self.estimate_pose_of_markers(corners, ids, marker_size)
                if should_filter:
                    self.observe(ids, detected_poses)
        if should_filter:
            camera_pose, marker_poses = self.get_poses()
            self._record_pose(camera_pose)
        else:
            _, marker_poses = self.get_poses()
            camera_pose = self.get_cam_estimate(iteration)
        return camera_pose, marker_poses, detected_poses

    def process_frame(self, frame, should_filter=True, iteration=0, marker_size: float = 0.16):
        """base_filter.py:173-212."""
        import cv2
        corners, ids, _ = tuple(self.detector.detectMarkers(frame))
        if ids is not None:
            frame = cv2.aruco.drawDetectedMarkers(frame, corners, ids)
        camera_pose, marker_poses, detected = self.process_detections(
            corners, ids, should_filter, iteration, marker_size)
        return frame, camera_pose, marker_poses, detected

    def save_map(self, filename: str) -> None:
        """base_filter.py:214-247, byte-compatible text format (outputs/map.txt)."""
        _, marker_poses = self.get_poses()
        index_to_id = {v: k for k, v in self.get_lm_estimates()}
        uncertainties = self.get_lm_uncertainties()
        with Path(filename).open("w", encoding="utf-8") as file:
            file.write("# landmark_id\n")
            file.write("# x y z\n")
            file.write("# uncertainty\n")
            file.write("\n")
            for i, pose in enumerate(marker_poses):
                file.write(f"{index_to_id[i]}\n")
                file.write(f"{', '.join(map(str, pose))}\n")
                file.write(f"{', '.join(map(str, uncertainties[i, :len(pose)]))}\n")
                file.write("\n")

    def load_map(self, filename: str) -> None:
        """base_filter.py:249-272 (dead code in the reference: ``self.filter`` does not
        exist there); here it adds each stored landmark with its diagonal."""
        with Path(filename).open("r", encoding="utf-8") as file:
            lines = file.readlines()[4:]
        for i in range(0, len(lines) - 2, 4):
            id_ = int(lines[i].strip())
            pose = np.array(lines[i + 1].strip().split(", "), np.float32)
            unc = np.array(lines[i + 2].strip().split(", "), np.float32)
            self.add_marker(id_, pose, unc)

    # pose history for run_offline's second pass (base_filter.py:369-381)
    def _record_pose(self, camera_pose) -> None:
        self._history.append(np.array(camera_pose, dtype=np.float64))

    def get_cam_estimate(self, iteration: int) -> np.ndarray:
        return self._history[iteration]

    def batch_optimize(self) -> None:
        """No-op: an EKF has nothing to re-optimise (main/run_offline.py:96)."""

    def observe(self, ids, poses) -> None:
        raise NotImplementedError(NOT_IMPLEMENTED_ERROR)

    def get_poses(self):
        raise NotImplementedError(NOT_IMPLEMENTED_ERROR)

    def get_lm_uncertainties(self):
        raise NotImplementedError(NOT_IMPLEMENTED_ERROR)

    def get_lm_estimates(self):
        raise NotImplementedError(NOT_IMPLEMENTED_ERROR)


class _CudaFilter(BaseFilter):
    KIND = _lib.KIND_EKF
    LM_DIMS = 3

    def __init__(self, initial_camera_pose, aruco_dict=None, *, camera_matrix=None,
                 dist_coeffs=None, quat_mode: str = "as_written", max_landmarks: int = 256,
                 noise: dict | None = None, device: int = 0, growable: bool = True) -> None:
        self._history = []
        self._h = None
        super().__init__(initial_camera_pose, None, aruco_dict, camera_matrix=camera_matrix,
                         dist_coeffs=dist_coeffs, device=device)
        modes = {"as_written": _lib.QUAT_AS_WRITTEN, "scalar_first": _lib.QUAT_SCALAR_FIRST}
        if quat_mode not in modes:
            raise ValueError(f"quat_mode must be one of {list(modes)}")
        self.quat_mode = quat_mode
        self._lib = _lib.lib()
        pose = _lib.as_f64(initial_camera_pose, (CAM_DIMS,))
        cfg = None
        if noise is not None:
            nz = _lib.Noise()
            _lib.check(self._lib.asb_default_noise(self.KIND, C.byref(nz)))
            for key, val in noise.items():
                setattr(nz, key, float(val))
            cfg = C.byref(nz)
        self._create_args = (modes[quat_mode], pose, cfg, int(device))
        self._noise_keep = cfg
        self._stream = None
        self.growable = bool(growable)
        self._h = self._create(int(max_landmarks))
        self.max_landmarks = int(max_landmarks)
        self.landmarks = {}

    def _create(self, max_landmarks: int):
        mode, pose, cfg, device = self._create_args
        h = C.c_void_p()
        _lib.check(self._lib.asb_filter_create(self.KIND, mode, _lib.dptr(pose), cfg, int(max_landmarks), device,
                                               C.byref(h)))
        return h

    def _grow(self, need: int) -> None:
        """The reference grows its state and covariance without bound (extended_kalman_filter.py:279-290); the
        device buffers are allocated at a capacity, so on overflow the filter moves to a handle of at least twice
        the capacity (state, covariance and id map copied, as the reference's re-allocation copies P)."""
        state, cov = self.state, self.uncertainty
        self._refresh_landmarks()
        ids = self._ids.copy()
        new_cap = max(2 * self.max_landmarks, int(need))
        new_h = self._create(new_cap)
        old, self._h = self._h, new_h
        self._lib.asb_filter_destroy(old)
        self.max_landmarks = new_cap
        if self._stream:
            _lib.check(self._lib.asb_filter_set_stream(self._h, C.c_void_p(self._stream)))
        if getattr(self, "_motion_window", 0):
            _lib.check(self._lib.asb_set_motion_model(self._h, self._motion_window))   # its history restarts
        if getattr(self, "_small_map_limit", None) is not None:
            _lib.check(self._lib.asb_set_small_map_limit(self._h, self._small_map_limit))
        if ids.size:
            self.load_checkpoint(state, cov, ids)
        else:
            self.state = state
            self.uncertainty = cov

    def _call_growing(self, fn, n_new_max: int):
        """Run a library call that may add landmarks; on ASB_ERR_CAPACITY (nothing was changed) grow and retry."""
        rc = fn()
        if rc == _lib.ERR_CAPACITY and self.growable:
            self._grow(self.num_landmarks + n_new_max)
            rc = fn()
        _lib.check(rc)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.asb_filter_destroy(h)
            except Exception:  # interpreter shutdown
                pass

    # ---- attributes the reference exposes (EKF:46-51) --------------------------------
    def _dims(self):
        nl, ns = C.c_int(), C.c_int()
        _lib.check(self._lib.asb_get_dims(self._h, C.byref(nl), C.byref(ns)))
        return nl.value, ns.value

    @property
    def num_landmarks(self) -> int:
        return self._dims()[0]

    def _refresh_landmarks(self) -> None:
        nl, _ = self._dims()
        if nl != len(self._id_order_cache()) or len(self.landmarks) > nl:
            ids = np.empty(max(nl, 1), dtype=np.int32)
            _lib.check(self._lib.asb_get_landmark_ids(self._h, _lib.iptr(ids), len(ids)))
            self._ids = ids[:nl].copy()
            # dict id -> index; a re-bound id keeps the LAST index, as dict assignment does.  Rebuilt only when the
            # map changed: at 2000 landmarks the rebuild costs ~200 us of Python per frame, twice the GPU's frame
            self.landmarks = {int(i): j for j, i in enumerate(self._ids)}

    def _id_order_cache(self):
        if not hasattr(self, "_ids"):
            self._ids = np.empty(0, dtype=np.int32)
        return self._ids

    @property
    def state(self) -> np.ndarray:
        _, n = self._dims()
        out = np.empty(n, dtype=np.float64)
        _lib.check(self._lib.asb_get_state(self._h, _lib.dptr(out), n))
        return out

    @state.setter
    def state(self, value) -> None:
        v = _lib.as_f64(value).reshape(-1)
        _lib.check(self._lib.asb_set_state(self._h, _lib.dptr(v), v.size))

    @property
    def uncertainty(self) -> np.ndarray:
        _, n = self._dims()
        out = np.empty((n, n), dtype=np.float64)
        _lib.check(self._lib.asb_get_cov(self._h, _lib.dptr(out), n))
        return out

    @uncertainty.setter
    def uncertainty(self, value) -> None:
        v = _lib.as_f64(value)
        _lib.check(self._lib.asb_set_cov(self._h, _lib.dptr(v), v.shape[0]))

    def set_landmarks(self, ids_in_index_order) -> None:
        """Replace the id->index map (checkpoint restore / oracle re-synchronisation)."""
        ids = _lib.as_i32(ids_in_index_order)
        _lib.check(self._lib.asb_set_landmarks(self._h, _lib.iptr(ids), ids.size))
        self._ids = ids.copy()
        self.landmarks = {int(i): j for j, i in enumerate(self._ids)}

    def load_checkpoint(self, state, uncertainty, ids_in_index_order) -> None:
        self.set_landmarks(ids_in_index_order)
        self.state = state
        self.uncertainty = uncertainty

    # ---- the per-frame API -----------------------------------------------------------
    def observe(self, ids, poses) -> None:
        """EKF.observe (EKF:58-82): add unseen ids in frame order, predict, update.
        Returns None, as the reference does."""
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        if ids.size != poses.shape[0]:
            raise ValueError(f"{ids.size} ids but {poses.shape[0]} poses")
        self._call_growing(lambda: self._lib.asb_observe(self._h, _lib.iptr(ids), _lib.dptr(poses), ids.size), ids.size)
        self._refresh_landmarks()

    def _observe_corners(self, ids, corners, marker_size: float) -> np.ndarray:
        """estimate_pose_of_markers + observe (BASE:196-204) as one library call; returns the
        detected poses (k,6) like ``estimate_pose_of_markers``."""
        from .ippe import _corner_array, _intrinsics
        ids = _lib.as_i32(ids)
        c = _corner_array(corners)
        km, dist = _intrinsics(self.calib_matrix, self.dist_coeffs)
        if ids.size != c.shape[0]:
            raise ValueError(f"{ids.size} ids but {c.shape[0]} corner sets")
        out = np.empty((c.shape[0], 6), dtype=np.float64)
        self._call_growing(lambda: self._lib.asb_observe_corners(
            self._h, _lib.iptr(ids), _lib.fptr(c), ids.size, _lib.dptr(km), _lib.dptr(dist), float(marker_size),
            _lib.dptr(out)), ids.size)
        self._refresh_landmarks()
        return out

    def correct_poses(self, ids, poses, alt_poses=None, threshold_deg: float = 50.0):
        """``BaseFilter.correct_poses`` (base_filter.py:274-325; QUAT_THRESHOLD = 50, :32): keep the observations
        whose orientation agrees with the landmark orientation in the state, drop unknown ids.  Returns
        ``(corrected_ids, corrected_poses)`` as the reference does.  ``alt_poses`` (k,6): the second IPPE solution
        of each observation (``estimate_both_poses``); it replaces a rejected first solution when it passes the
        same test (extension, README "orientation ambiguity")."""
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        if ids.size != poses.shape[0]:
            raise ValueError("ids and poses disagree on the number of observations")
        alt = None if alt_poses is None else _lib.as_f64(alt_poses, (-1, 6))
        if alt is not None and alt.shape != poses.shape:
            raise ValueError("alt_poses must have the shape of poses")
        choice = np.zeros(ids.size, dtype=np.int32)
        _lib.check(self._lib.asb_correct_poses(self._h, _lib.iptr(ids), _lib.dptr(poses),
                                               None if alt is None else _lib.dptr(alt), ids.size,
                                               float(threshold_deg), _lib.iptr(choice)))
        out_ids = [int(i) for i, c in zip(ids, choice) if c]
        out_poses = [poses[o] if c == 1 else alt[o] for o, c in enumerate(choice) if c]
        return out_ids, out_poses

    def set_motion_model(self, window: int = 10) -> None:
        """Opt-in moving-average motion model of the reference's README (README.md:55-56; MOVING_AVG_WINDOW = 10,
        extended_kalman_filter.py:19, never read there): before each ``observe`` the camera position moves by
        ``(X_k-1 - X_k-n) / n``.  ``window=0`` (default) is the reference code's behaviour: predict leaves the state."""
        _lib.check(self._lib.asb_set_motion_model(self._h, int(window)))
        self._motion_window = int(window)

    def set_small_map_limit(self, n_state_max: int) -> None:
        """``observe`` on a map of at most ``n_state_max`` state dimensions runs as one kernel launch per frame
        (default 192); larger maps, or ``0``, use the pipelined multi-SM engine."""
        _lib.check(self._lib.asb_set_small_map_limit(self._h, int(n_state_max)))
        self._small_map_limit = int(n_state_max)

    def predict(self) -> None:
        _lib.check(self._lib.asb_predict(self._h))

    def update(self, ids, poses) -> None:
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        if ids.size != poses.shape[0]:
            raise ValueError(f"{ids.size} ids but {poses.shape[0]} poses")
        _lib.check(self._lib.asb_update(self._h, _lib.iptr(ids), _lib.dptr(poses), ids.size))

    def add_marker(self, idx, pose, uncertainity=None) -> None:
        pose = _lib.as_f64(pose).reshape(-1)
        if pose.size < 6:
            pose = np.concatenate([pose, np.zeros(6 - pose.size)])
        unc = None
        if uncertainity is not None:
            u = np.asarray(uncertainity, dtype=np.float64)
            if u.ndim == 2:
                u = np.diagonal(u)  # eye * M keeps only M's diagonal (EKF:280-281)
            u = np.ascontiguousarray(np.broadcast_to(u, (self.LM_DIMS,)))
            unc = _lib.dptr(u)
        self._call_growing(lambda: self._lib.asb_add_marker(self._h, int(idx), _lib.dptr(pose), unc), 1)
        self._refresh_landmarks()

    def get_poses(self):
        """(camera_pose (10,), marker_poses (n, d)) — copies, not views (EKF:84-89)."""
        s = self.state
        return s[:CAM_DIMS], s[CAM_DIMS:].reshape(-1, self.LM_DIMS)

    def get_lm_uncertainties(self) -> np.ndarray:
        _, n = self._dims()
        out = np.empty(n, dtype=np.float64)
        _lib.check(self._lib.asb_get_cov_diag(self._h, _lib.dptr(out), n))
        return out[CAM_DIMS:].reshape(-1, self.LM_DIMS)

    def get_lm_estimates(self):
        self._refresh_landmarks()
        return self.landmarks.items()

    # ---- extras used by tests / bench ------------------------------------------------
    def synchronize(self) -> None:
        _lib.check(self._lib.asb_filter_synchronize(self._h))

    def set_stream(self, cuda_stream: int | None) -> None:
        self._stream = cuda_stream or None
        _lib.check(self._lib.asb_filter_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def stream_upload(self, ids, poses, offsets) -> None:
        ids = _lib.as_i32(ids)
        poses = _lib.as_f64(poses, (-1, 6))
        off = np.ascontiguousarray(np.asarray(offsets, dtype=np.int64))
        _lib.check(self._lib.asb_stream_upload(self._h, _lib.iptr(ids), _lib.dptr(poses),
                                               _lib.lptr(off), off.size - 1))

    def stream_run(self, t0: int, t1: int) -> None:
        _lib.check(self._lib.asb_stream_run(self._h, int(t0), int(t1)))
        self._refresh_landmarks()

    def last_update(self):
        """(nu, blocks (k, r, 10+d), dx) of the last update, for parity tests."""
        _, n = self._dims()
        m = C.c_int()
        _lib.check(self._lib.asb_get_last_update(self._h, None, None, None, C.byref(m)))
        r = 3 if self.KIND == _lib.KIND_EKF else 7
        k = m.value // r
        nu = np.empty(m.value)
        blocks = np.empty((k, r, CAM_DIMS + self.LM_DIMS))
        dx = np.empty(n)
        _lib.check(self._lib.asb_get_last_update(self._h, _lib.dptr(nu), _lib.dptr(blocks),
                                                 _lib.dptr(dx), C.byref(m)))
        return nu, blocks, dx

    def device_views(self):
        ds, dc, ld = C.c_void_p(), C.c_void_p(), C.c_int()
        _lib.check(self._lib.asb_device_views(self._h, C.byref(ds), C.byref(dc), C.byref(ld)))
        return ds.value, dc.value, ld.value


class EKF(_CudaFilter):
    """``EKF(initial_camera_pose, aruco_dict=None)`` (extended_kalman_filter.py:37-56):
    camera 10 dims + 3 per landmark, 3 measurement rows per marker."""
    KIND = _lib.KIND_EKF
    LM_DIMS = 3


class EKF_Rotations(_CudaFilter):  # noqa: N801 — the reference's class name
    """``EKF_Rotations(initial_camera_pose)`` (ekf_with_rotations.py:43-64): 10 dims per
    landmark (xyz, quaternion, error), 7 measurement rows per marker."""
    KIND = _lib.KIND_ROT
    LM_DIMS = 10

    def __init__(self, initial_camera_pose, **kw) -> None:
        kw.pop("quat_mode", None)
        super().__init__(initial_camera_pose, None, quat_mode="scalar_first", **kw)
