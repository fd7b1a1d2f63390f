"""numpy restatement of the scipy ``Rotation`` / sympy ``Quaternion`` arithmetic
the reference filters call.

TEST INFRASTRUCTURE (CPU oracle) — only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import this.

Third-party arithmetic restated here (absent from /root/reference, pinned in
/root/reference/requirements.txt): scipy==1.14.1 ``Rotation`` (installed here:
1.18.1) and sympy==1.13.3 ``Quaternion`` (installed: 1.14.0).  Conventions are
pinned by ``tests/test_oracle_rotations.py`` against the installed scipy and by
the golden vectors produced by the reference itself.

All quaternions in this module are scalar-FIRST ``[w, x, y, z]`` unless a
function name says otherwise.
"""
from __future__ import annotations

import numpy as np


def quat_normalize(q):
    q = np.asarray(q, dtype=np.float64)
    return q / np.sqrt(q @ q)


def hamilton(p, q):
    """p ⊗ q, scalar-first (sympy ``Quaternion.mul``; scipy ``compose_quat``)."""
    p0, p1, p2, p3 = p
    q0, q1, q2, q3 = q
    return np.array([
        p0 * q0 - p1 * q1 - p2 * q2 - p3 * q3,
        p0 * q1 + p1 * q0 + p2 * q3 - p3 * q2,
        p0 * q2 - p1 * q3 + p2 * q0 + p3 * q1,
        p0 * q3 + p1 * q2 - p2 * q1 + p3 * q0,
    ])


def left_matrix(p):
    """L(p) with p ⊗ q = L(p) q."""
    p0, p1, p2, p3 = p
    return np.array([[p0, -p1, -p2, -p3],
                     [p1, p0, -p3, p2],
                     [p2, p3, p0, -p1],
                     [p3, -p2, p1, p0]])


def right_matrix(q):
    """Rm(q) with p ⊗ q = Rm(q) p."""
    q0, q1, q2, q3 = q
    return np.array([[q0, -q1, -q2, -q3],
                     [q1, q0, q3, -q2],
                     [q2, -q3, q0, q1],
                     [q3, q2, -q1, q0]])


def quad_form_matrix(u):
    """M(u): the homogeneous rotation quadratic form of sympy
    ``Quaternion.to_rotation_matrix`` before its 1/|u|^2 scale
    (extended_kalman_filter.py:317)."""
    a, b, c, d = u
    return np.array([
        [a * a + b * b - c * c - d * d, 2 * (b * c - d * a), 2 * (b * d + c * a)],
        [2 * (b * c + d * a), a * a - b * b + c * c - d * d, 2 * (c * d - b * a)],
        [2 * (b * d - c * a), 2 * (c * d + b * a), a * a - b * b - c * c + d * d],
    ])


def quad_form_derivs(u):
    """dM/du_k, k = 0..3, each 3x3."""
    a, b, c, d = u
    return 2.0 * np.array([
        [[a, -d, c], [d, a, -b], [-c, b, a]],
        [[b, c, d], [c, -b, -a], [d, a, -b]],
        [[-c, b, a], [b, c, d], [-a, d, -c]],
        [[-d, -a, b], [a, -d, c], [b, c, d]],
    ])


def rotation_matrix(q):
    """scipy ``Rotation.from_quat(q, scalar_first=True).as_matrix()``: q is
    normalised first, then the unit-quaternion matrix."""
    return quad_form_matrix(quat_normalize(q))


def compose_normalized(dq, q):
    """``(Rotation.from_quat(dq) * Rotation.from_quat(q)).as_quat()`` with both
    inputs scalar-first: normalise each, Hamilton product, normalise."""
    return quat_normalize(hamilton(quat_normalize(dq), quat_normalize(q)))


def euler_xyz_quat(angles):
    """``Rotation.from_euler('xyz', angles).as_quat(scalar_first=True)``:
    extrinsic x, y, z  =>  q = qz ⊗ qy ⊗ qx, no sign canonicalisation
    (ekf_with_rotations.py:216-219)."""
    a0, a1, a2 = (0.5 * float(v) for v in angles)
    qx = np.array([np.cos(a0), np.sin(a0), 0.0, 0.0])
    qy = np.array([np.cos(a1), 0.0, np.sin(a1), 0.0])
    qz = np.array([np.cos(a2), 0.0, 0.0, np.sin(a2)])
    return quat_normalize(hamilton(qz, hamilton(qy, qx)))


def euler_xyz_matrix(angles):
    """``Rotation.from_euler('xyz', angles).as_matrix()`` (ekf_with_rotations.py:307-310)."""
    return quad_form_matrix(euler_xyz_quat(angles))


def matrix_to_quat(m):
    """``Rotation.from_matrix(m).as_quat(scalar_first=True)`` for an (already
    orthogonal to 1e-12) matrix: scipy's largest-of-(m00, m11, m22, trace) branch,
    first maximum wins, result normalised, sign NOT canonicalised
    (ekf_with_rotations.py:314-315)."""
    m = np.asarray(m, dtype=np.float64)
    tr = m[0, 0] + m[1, 1] + m[2, 2]
    choice = int(np.argmax([m[0, 0], m[1, 1], m[2, 2], tr]))
    if choice == 3:
        x, y, z, w = m[2, 1] - m[1, 2], m[0, 2] - m[2, 0], m[1, 0] - m[0, 1], 1 + tr
    elif choice == 0:
        x, y, z, w = 1 - tr + 2 * m[0, 0], m[1, 0] + m[0, 1], m[2, 0] + m[0, 2], m[2, 1] - m[1, 2]
    elif choice == 1:
        x, y, z, w = m[1, 0] + m[0, 1], 1 - tr + 2 * m[1, 1], m[2, 1] + m[1, 2], m[0, 2] - m[2, 0]
    else:
        x, y, z, w = m[2, 0] + m[0, 2], m[2, 1] + m[1, 2], 1 - tr + 2 * m[2, 2], m[1, 0] - m[0, 1]
    return quat_normalize(np.array([w, x, y, z]))
