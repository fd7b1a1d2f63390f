"""ctypes binding of libarucoslam_b200.so (C ABI: include/aruco_slam_b200.h).

The library is the product; this module only declares its entry points.  There is
no Python / numpy / CPU implementation of the hot path behind it: if the shared
library is missing or no CUDA device is present, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
# ASB_LIB_PATH: A/B experiments against another build of the same library (tools/); never a fallback
LIB_PATH = Path(os.environ.get("ASB_LIB_PATH") or PKG_DIR / "libarucoslam_b200.so")
CSRC = PKG_DIR / "csrc"

ASB_OK = 0
ERR_BAD_HANDLE, ERR_BAD_ARG, ERR_UNKNOWN_ID, ERR_CAPACITY, ERR_NOT_SPD, ERR_CUDA, ERR_NO_STREAM = \
    -1, -2, -3, -4, -5, -6, -7
KIND_EKF, KIND_ROT = 0, 1
QUAT_AS_WRITTEN, QUAT_SCALAR_FIRST = 0, 1


class AsbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[asb {code}] {msg}")
        self.code = code


class Noise(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("init_cam", "init_lm", "r", "q_cam", "q_err", "q_lm")]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)
_fp = C.POINTER(C.c_float)
_vp = C.c_void_p
_intp = C.POINTER(C.c_int)

# name -> (restype, argtypes); every symbol include/aruco_slam_b200.h declares
SIGNATURES = {
    "asb_abi_version": (C.c_int, []),
    "asb_last_error": (C.c_char_p, []),
    "asb_device_count": (C.c_int, [_intp]),
    "asb_default_noise": (C.c_int, [C.c_int, C.POINTER(Noise)]),
    "asb_filter_create": (C.c_int, [C.c_int, C.c_int, _dp, C.POINTER(Noise), C.c_int, C.c_int,
                                    C.POINTER(_vp)]),
    "asb_filter_destroy": (C.c_int, [_vp]),
    "asb_filter_set_stream": (C.c_int, [_vp, _vp]),
    "asb_filter_synchronize": (C.c_int, [_vp]),
    "asb_observe": (C.c_int, [_vp, _ip, _dp, C.c_int]),
    "asb_observe_corners": (C.c_int, [_vp, _ip, _fp, C.c_int, _dp, _dp, C.c_double, _dp]),
    "asb_predict": (C.c_int, [_vp]),
    "asb_update": (C.c_int, [_vp, _ip, _dp, C.c_int]),
    "asb_add_marker": (C.c_int, [_vp, C.c_int32, _dp, _dp]),
    "asb_stream_upload": (C.c_int, [_vp, _ip, _dp, _lp, C.c_int]),
    "asb_stream_run": (C.c_int, [_vp, C.c_int, C.c_int]),
    "asb_get_dims": (C.c_int, [_vp, _intp, _intp]),
    "asb_get_state": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_set_state": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_get_cov": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_set_cov": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_get_cov_diag": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_get_landmark_ids": (C.c_int, [_vp, _ip, C.c_int]),
    "asb_set_landmarks": (C.c_int, [_vp, _ip, C.c_int]),
    "asb_device_views": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), _intp]),
    "asb_set_motion_model": (C.c_int, [_vp, C.c_int]),
    "asb_set_small_map_limit": (C.c_int, [_vp, C.c_int]),
    "asb_correct_poses": (C.c_int, [_vp, _ip, _dp, _dp, C.c_int, C.c_double, _ip]),
    "asb_get_last_update": (C.c_int, [_vp, _dp, _dp, _dp, _intp]),
    "asb_batch_create": (C.c_int, [C.c_int, C.c_int, _dp, C.POINTER(Noise), C.c_int, C.c_int,
                                   C.c_int, C.c_int, C.POINTER(_vp)]),
    "asb_batch_destroy": (C.c_int, [_vp]),
    "asb_batch_reset": (C.c_int, [_vp]),
    "asb_batch_set_engine": (C.c_int, [_vp, C.c_int]),
    "asb_batch_get_engine": (C.c_int, [_vp, _intp]),
    "asb_batch_set_stream": (C.c_int, [_vp, _vp]),
    "asb_batch_synchronize": (C.c_int, [_vp]),
    "asb_batch_upload": (C.c_int, [_vp, _ip, _dp, _lp, C.c_int]),
    "asb_batch_upload_tiled": (C.c_int, [_vp, _ip, _dp, _lp, C.c_int, C.c_int]),
    "asb_batch_run": (C.c_int, [_vp, C.c_int, C.c_int]),
    "asb_batch_get_dims": (C.c_int, [_vp, C.c_int, _intp, _intp]),
    "asb_batch_get_state": (C.c_int, [_vp, C.c_int, _dp, C.c_int]),
    "asb_batch_get_cov": (C.c_int, [_vp, C.c_int, _dp, C.c_int]),
    "asb_batch_get_landmark_ids": (C.c_int, [_vp, C.c_int, _ip, C.c_int]),
    "asb_batch_set_landmarks": (C.c_int, [_vp, C.c_int, _ip, C.c_int]),
    "asb_batch_set_state": (C.c_int, [_vp, C.c_int, _dp, C.c_int]),
    "asb_batch_set_cov": (C.c_int, [_vp, C.c_int, _dp, C.c_int]),
    "asb_batch_get_cameras": (C.c_int, [_vp, _dp, _ip]),
    "asb_batch_device_views": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), _intp, _intp]),
    "asb_batch_profile": (C.c_int, [_vp, _dp]),
    "asb_shard_partition": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _intp, _intp, _intp]),
    "asb_shard_unique_id": (C.c_int, [_vp, C.c_int]),
    "asb_shard_create": (C.c_int, [C.c_int, C.c_int, _dp, C.POINTER(Noise), C.c_int, C.c_int, C.c_int, C.c_int,
                                   _vp, C.POINTER(_vp)]),
    "asb_shard_destroy": (C.c_int, [_vp]),
    "asb_shard_set_stream": (C.c_int, [_vp, _vp]),
    "asb_shard_synchronize": (C.c_int, [_vp]),
    "asb_shard_observe": (C.c_int, [_vp, _ip, _dp, C.c_int]),
    "asb_shard_stream_upload": (C.c_int, [_vp, _ip, _dp, _lp, C.c_int]),
    "asb_shard_stream_run": (C.c_int, [_vp, C.c_int, C.c_int]),
    "asb_shard_get_dims": (C.c_int, [_vp, _intp, _intp]),
    "asb_shard_get_state": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_shard_get_cov_rows": (C.c_int, [_vp, _dp, C.c_longlong, _intp]),
    "asb_shard_get_landmark_ids": (C.c_int, [_vp, _ip, C.c_int]),
    "asb_ippe_square": (C.c_int, [_fp, C.c_int64, _dp, _dp, C.c_double, _dp, C.c_int]),
    "asb_ippe_square_both": (C.c_int, [_fp, C.c_int64, _dp, _dp, C.c_double, _dp, _dp, _fp, C.c_int]),
    "asb_ippe_square_static": (C.c_int, [_fp, C.c_int64, _dp, _dp, C.c_double, _dp, C.c_int]),
    "asb_ippe_square_device": (C.c_int, [_vp, C.c_int64, _dp, _dp, C.c_double, _vp, _vp, _vp, _vp]),
    "asb_bench_sweep": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, _dp]),
    "asb_front_profile": (C.c_int, [_vp, _dp]),
    "asb_timeline": (C.c_int, [_vp, _dp]),
    "asb_cta_profile": (C.c_int, [_vp, _dp, C.c_int]),
    "asb_latency_probe": (C.c_int, [C.c_int, _dp]),
    "asb_stream_bandwidth": (C.c_int, [C.c_int, C.c_int64, C.c_int, _dp]),
    "asb_tilepattern_bandwidth": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _dp]),
    "asb_fp64_peak": (C.c_int, [C.c_int, C.c_int, _dp]),
    "asb_copy_bandwidth": (C.c_int, [C.c_int, C.c_int64, _dp]),
}

_lib = None


def build(verbose: bool = False) -> Path:
    """Compile the CUDA sources for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", str(CSRC), "-j", str(os.cpu_count() or 4)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libarucoslam_b200.so failed")
    return LIB_PATH


def lib() -> C.CDLL:
    """Load the shared library (fails loudly; no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"or `make -C {CSRC}`.  aruco_slam_b200 has no CPU fallback.")
    handle = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(handle, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if handle.asb_abi_version() != 1:
        raise ImportError("libarucoslam_b200.so ABI version mismatch")
    _lib = handle
    return _lib


def check(code: int) -> None:
    if code == ASB_OK:
        return
    msg = lib().asb_last_error().decode("utf-8", "replace")
    if code == ERR_UNKNOWN_ID:
        raise KeyError(msg)  # the reference raises KeyError (extended_kalman_filter.py:180)
    raise AsbError(code, msg)


def dptr(a: np.ndarray):
    return a.ctypes.data_as(_dp)


def iptr(a: np.ndarray):
    return a.ctypes.data_as(_ip)


def lptr(a: np.ndarray):
    return a.ctypes.data_as(_lp)


def fptr(a: np.ndarray):
    return a.ctypes.data_as(_fp)


def device_count() -> int:
    n = C.c_int(0)
    rc = lib().asb_device_count(C.byref(n))
    return n.value if rc == ASB_OK else 0


def as_f64(a, shape=None) -> np.ndarray:
    out = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    if shape is not None:
        out = out.reshape(shape)
    return out


def as_i32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a).astype(np.int32, copy=False).reshape(-1))
