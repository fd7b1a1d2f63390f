"""Batched square-marker pose estimation (replaces the per-marker
``cv2.solvePnP(..., flags=cv2.SOLVEPNP_IPPE_SQUARE)`` loop of
``BaseFilter.estimate_pose_of_markers``, filters/base_filter.py:92-171)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _corner_array(corners) -> np.ndarray:
    """Accept what ``detectMarkers`` returns (tuple of (1,4,2) float32) or any
    array reshapeable to (k,4,2)."""
    if isinstance(corners, (tuple, list)):
        if len(corners) == 0:
            return np.zeros((0, 4, 2), dtype=np.float32)
        corners = np.stack([np.asarray(c, dtype=np.float32).reshape(4, 2) for c in corners])
    return np.ascontiguousarray(np.asarray(corners, dtype=np.float32).reshape(-1, 4, 2))


def _intrinsics(camera_matrix, dist_coeffs):
    km = _lib.as_f64(camera_matrix, (9,))
    dist = np.zeros(5)
    d = _lib.as_f64(dist_coeffs).reshape(-1)
    if d.size > 5 and np.any(d[5:] != 0.0):
        # cv2.solvePnP honours the 8 / 12 / 14-coefficient models; only k1 k2 p1 p2 k3 is implemented here
        raise ValueError("only the 5-coefficient distortion model (k1, k2, p1, p2, k3) is implemented; "
                         f"got {d.size} coefficients with non-zero higher-order terms")
    dist[:min(5, d.size)] = d[:5]
    return km, dist


def estimate_pose_of_markers(corners, camera_matrix, dist_coeffs, marker_size: float = 0.16,
                             device: int = 0, out: np.ndarray | None = None) -> np.ndarray:
    """(k,4,2) float32 corners -> (k,6) float64 ``[tvec, rvec]`` rows (host in, host out).
    ``out``: optional (k,6) float64 C-contiguous result array (e.g. a view of pinned memory: with
    page-locked ``corners`` / ``out`` the two copies run at full PCIe speed)."""
    c = _corner_array(corners)
    km, dist = _intrinsics(camera_matrix, dist_coeffs)
    if out is None:
        out = np.empty((c.shape[0], 6), dtype=np.float64)
    elif out.shape != (c.shape[0], 6) or out.dtype != np.float64 or not out.flags.c_contiguous:
        raise ValueError("out must be a C-contiguous (k, 6) float64 array")
    _lib.check(_lib.lib().asb_ippe_square(_lib.fptr(c), c.shape[0], _lib.dptr(km), _lib.dptr(dist),
                                          float(marker_size), _lib.dptr(out), int(device)))
    return out


def estimate_both_poses(corners, camera_matrix, dist_coeffs, marker_size: float = 0.16, device: int = 0):
    """``cv2.solvePnPGeneric(..., SOLVEPNP_IPPE_SQUARE)`` for every corner set: (poses (k,2,6) float64
    ``[tvec, rvec]`` with the better solution first, errors (k,2) float32)."""
    c = _corner_array(corners)
    km, dist = _intrinsics(camera_matrix, dist_coeffs)
    n = c.shape[0]
    first, second = np.empty((n, 6)), np.empty((n, 6))
    errs = np.empty((n, 2), dtype=np.float32)
    _lib.check(_lib.lib().asb_ippe_square_both(_lib.fptr(c), n, _lib.dptr(km), _lib.dptr(dist), float(marker_size),
                                               _lib.dptr(first), _lib.dptr(second), _lib.fptr(errs), int(device)))
    return np.stack([first, second], axis=1), errs


def estimate_pose_device(d_corners: int, n: int, camera_matrix, dist_coeffs, marker_size: float,
                         d_out: int, d_out2: int | None = None, d_err: int | None = None,
                         stream: int | None = None) -> None:
    """Device-pointer variant (raw addresses, e.g. ``tensor.data_ptr()``), asynchronous."""
    km, dist = _intrinsics(camera_matrix, dist_coeffs)
    _lib.check(_lib.lib().asb_ippe_square_device(
        C.c_void_p(d_corners), int(n), _lib.dptr(km), _lib.dptr(dist), float(marker_size),
        C.c_void_p(d_out), C.c_void_p(d_out2 or 0), C.c_void_p(d_err or 0), C.c_void_p(stream or 0)))
