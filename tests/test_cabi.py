"""CPU: the C-ABI shared library builds/loads and exports every symbol that
include/aruco_slam_b200.h declares; without a GPU the product path fails loudly
(no CPU fallback, no routing through oracle/)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "aruco_slam_b200.h"


@pytest.fixture(scope="module")
def lib():
    from aruco_slam_b200 import _lib
    if not _lib.LIB_PATH.exists():
        _lib.build()
    return _lib.lib()


def declared_symbols():
    text = HEADER.read_text()
    return sorted(set(re.findall(r"ASB_API\s+(?:const\s+char\*|int)\s+(asb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    from aruco_slam_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 40
    raw = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in names:
        assert hasattr(raw, name), f"{name} declared in the header but not exported"
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert lib.asb_abi_version() == 1


def test_default_noise_matches_reference_constants(lib):
    from aruco_slam_b200 import _lib
    nz = _lib.Noise()
    assert lib.asb_default_noise(_lib.KIND_EKF, ctypes.byref(nz)) == 0
    assert (nz.init_cam, nz.init_lm, nz.r, nz.q_cam, nz.q_err, nz.q_lm) == (0.1, 0.7, 0.9, 0.3, 0.5, 0.01)
    assert lib.asb_default_noise(_lib.KIND_ROT, ctypes.byref(nz)) == 0
    assert nz.q_cam == 0.2
    assert lib.asb_default_noise(7, ctypes.byref(nz)) == _lib.ERR_BAD_ARG
    assert b"bad kind" in lib.asb_last_error()


def test_no_gpu_fails_loudly(lib):
    from aruco_slam_b200 import _lib
    if _lib.device_count() > 0:
        pytest.skip("a GPU is present")
    import numpy as np
    from aruco_slam_b200 import EKF, FilterBatch, estimate_pose_of_markers
    eye = np.eye(3)
    with pytest.raises(_lib.AsbError):
        EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=eye, dist_coeffs=np.zeros(5))
    with pytest.raises(_lib.AsbError):
        FilterBatch("ekf", 2, 4, 4)
    with pytest.raises(_lib.AsbError):
        estimate_pose_of_markers(np.ones((1, 4, 2), np.float32), eye, np.zeros(5))


def test_product_never_imports_oracle():
    for path in (ROOT / "aruco_slam_b200").rglob("*.py"):
        text = path.read_text()
        assert "oracle" not in text.replace("oracle re-synchronisation", ""), path
