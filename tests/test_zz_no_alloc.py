"""GPU: the allocation-free entry point asb_ippe_square_static (static __device__ scratch, no
cudaMalloc) against the existing asb_ippe_square, through the C ABI, on two calls in a row.

Collected last on purpose.  Same kernel, same inputs: the
results must be bit-identical (tolerance 0; the IPPE parity tests against cv2 use 1e-9)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX = 16384  # ASB_IPPE_STATIC_MAX in include/aruco_slam_b200.h


def _call(fn, corners, km, dist, size=0.16):
    from aruco_slam_b200 import _lib
    out = np.full((corners.shape[0], 6), np.nan)
    rc = fn(_lib.fptr(corners), corners.shape[0], _lib.dptr(km), _lib.dptr(dist), float(size), _lib.dptr(out), 0)
    return rc, out


def _intrinsics(calibration):
    from aruco_slam_b200 import _lib
    km, dist = calibration
    return _lib.as_f64(km, (9,)), _lib.as_f64(dist).reshape(-1)[:5].copy()


def test_static_scratch_matches_allocating_entry_point(calibration):
    from aruco_slam_b200 import _lib, synth
    if _lib.device_count() == 0:
        pytest.skip("no CUDA device")
    lib = _lib.lib()
    km, dist = _intrinsics(calibration)
    # calls in a row on the shared scratch: the second is smaller and different, so stale rows
    # would show; then the maximum size
    for seed, n in ((11, 5000), (12, 1237), (13, STATIC_MAX)):
        corners = np.ascontiguousarray(synth.make_corner_sets(n, km.reshape(3, 3), dist, seed=seed)[:n])
        assert corners.shape == (n, 4, 2) and corners.dtype == np.float32
        rc1, ref = _call(lib.asb_ippe_square, corners, km, dist)
        rc2, got = _call(lib.asb_ippe_square_static, corners, km, dist)
        assert rc1 == 0 and rc2 == 0, lib.asb_last_error()
        assert np.isfinite(ref).all()
        assert np.array_equal(got, ref), float(np.nanmax(np.abs(got - ref)))
    text = (_lib.PKG_DIR.parent / "include" / "aruco_slam_b200.h").read_text()
    assert f"#define ASB_IPPE_STATIC_MAX {STATIC_MAX}" in text


def test_static_scratch_rejects_oversize(calibration):
    from aruco_slam_b200 import _lib
    if _lib.device_count() == 0:
        pytest.skip("no CUDA device")
    lib = _lib.lib()
    km, dist = _intrinsics(calibration)
    corners = np.ones((STATIC_MAX + 1, 4, 2), dtype=np.float32)
    rc, _ = _call(lib.asb_ippe_square_static, corners, km, dist)
    assert rc == _lib.ERR_CAPACITY
    assert b"ASB_IPPE_STATIC_MAX" in lib.asb_last_error()
    # n = 0 is a no-op like the allocating entry point
    rc, _ = _call(lib.asb_ippe_square_static, corners[:0], km, dist)
    assert rc == 0
