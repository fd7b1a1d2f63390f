"""CPU: oracle/rotations.py against the installed scipy Rotation (third-party
arithmetic the reference calls)."""
import numpy as np
import pytest

from oracle import rotations as rot

sp = pytest.importorskip("scipy.spatial.transform")
R = sp.Rotation


def test_euler_quat_sign_and_value():
    rng = np.random.default_rng(0)
    for _ in range(200):
        a = rng.uniform(-3.5, 3.5, 3)
        ref = R.from_euler("xyz", a).as_quat(scalar_first=True)
        assert np.allclose(rot.euler_xyz_quat(a), ref, atol=1e-15, rtol=0)
        assert np.allclose(rot.euler_xyz_matrix(a), R.from_euler("xyz", a).as_matrix(), atol=1e-15)


def test_matrix_to_quat_branches():
    rng = np.random.default_rng(1)
    for _ in range(500):
        q = rng.normal(size=4)
        m = R.from_quat(q, scalar_first=True).as_matrix()
        ref = R.from_matrix(m).as_quat(scalar_first=True)
        assert np.allclose(rot.matrix_to_quat(m), ref, atol=1e-14, rtol=0)


def test_compose_and_matrix():
    rng = np.random.default_rng(2)
    for _ in range(200):
        p, q = rng.normal(size=4), rng.normal(size=4)
        ref = (R.from_quat(p, scalar_first=True) * R.from_quat(q, scalar_first=True)) \
            .as_quat(scalar_first=True)
        assert np.allclose(rot.compose_normalized(p, q), ref, atol=1e-15, rtol=0)
        assert np.allclose(rot.rotation_matrix(q), R.from_quat(q, scalar_first=True).as_matrix(),
                           atol=1e-15)
        assert np.allclose(rot.left_matrix(p) @ q, rot.hamilton(p, q))
        assert np.allclose(rot.right_matrix(q) @ p, rot.hamilton(p, q))
