"""CPU oracle: numpy restatement of ``cv2.solvePnP(..., SOLVEPNP_IPPE_SQUARE)``
as called by the reference (/root/reference/filters/base_filter.py:156-164).

TEST INFRASTRUCTURE — not the product (see oracle/ekf_oracle.py header).

The arithmetic lives in a third-party dependency that is ABSENT from
/root/reference: ``opencv_python==4.10.0.84`` (requirements.txt:4; installed in
this image: 4.13.0).  What is restated is the published IPPE algorithm
(Collins & Bartoli, "Infinitesimal Plane-based Pose Estimation", IJCV 2014) in
the form OpenCV's ``calib3d`` uses for square targets, plus
``undistortPoints``' fixed 5-iteration inverse of the 5-coefficient
Brown-Conrady model.  PARITY PINNING: the reference holds no test for this
path; the restatement is pinned against the installed ``cv2`` called exactly as
the reference calls it (tests/golden/ippe_golden.npz, made by
oracle/make_golden.py; live cross-check in tests/test_oracle_ippe.py).

Vectorised over B corner sets: corners (B,4,2) float32 -> (B,6) float64
``[tvec, rvec]`` rows as ``estimate_pose_of_markers`` stacks them
(base_filter.py:165-171).
"""
from __future__ import annotations

import numpy as np


def undistort_normalize(corners, camera_matrix, dist_coeffs):
    """``cv2.undistortPoints`` with default criteria (5 iterations), k1 k2 p1 p2 k3.
    Input float32 pixels -> float64 math -> result ROUNDED TO float32 (the output
    array takes the input's depth), returned as float64 values."""
    k = np.zeros(5)
    d = np.asarray(dist_coeffs, dtype=np.float64).ravel()
    k[:min(5, d.size)] = d[:5]
    k1, k2, p1, p2, k3 = k
    km = np.asarray(camera_matrix, dtype=np.float64)
    fx, fy, cx, cy = km[0, 0], km[1, 1], km[0, 2], km[1, 2]
    ifx, ify = 1.0 / fx, 1.0 / fy
    u = corners[..., 0].astype(np.float64)
    v = corners[..., 1].astype(np.float64)
    x0 = (u - cx) * ifx
    y0 = (v - cy) * ify
    x, y = x0.copy(), y0.copy()
    frozen = np.zeros(x.shape, dtype=bool)
    for _ in range(5):
        r2 = x * x + y * y
        icdist = 1.0 / (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2)
        neg = (icdist < 0) & ~frozen
        dx = 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
        dy = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y
        xn = (x0 - dx) * icdist
        yn = (y0 - dy) * icdist
        # icdist < 0: OpenCV falls back to the undistorted-less start value and stops
        x = np.where(frozen, x, np.where(neg, x0, xn))
        y = np.where(frozen, y, np.where(neg, y0, yn))
        frozen |= neg
    return x.astype(np.float32).astype(np.float64), y.astype(np.float32).astype(np.float64)


def _homography_square(x, y, half):
    """Exact homography (H22 = 1) taking the canonical square
    (-h,+h), (+h,+h), (+h,-h), (-h,-h) to the 4 normalised points.
    Projective square->quad map on the unit square, then the affine pre-scale."""
    # unit square (0,0),(1,0),(1,1),(0,1)  <-  corners 3, 2, 1, 0
    x0, x1, x2, x3 = x[:, 3], x[:, 2], x[:, 1], x[:, 0]
    y0, y1, y2, y3 = y[:, 3], y[:, 2], y[:, 1], y[:, 0]
    sx = x0 - x1 + x2 - x3
    sy = y0 - y1 + y2 - y3
    dx1, dx2 = x1 - x2, x3 - x2
    dy1, dy2 = y1 - y2, y3 - y2
    den = dx1 * dy2 - dy1 * dx2
    g = (sx * dy2 - sy * dx2) / den
    hh = (dx1 * sy - dy1 * sx) / den
    a = x1 - x0 + g * x1
    b = x3 - x0 + hh * x3
    c = x0
    d = y1 - y0 + g * y1
    e = y3 - y0 + hh * y3
    f = y0
    # (X,Y) -> s = X/(2h) + 1/2, t = Y/(2h) + 1/2
    i2h = 1.0 / (2.0 * half)
    h22 = 0.5 * (g + hh) + 1.0
    hm = np.empty((x.shape[0], 3, 3))
    hm[:, 0, 0] = a * i2h / h22
    hm[:, 0, 1] = b * i2h / h22
    hm[:, 0, 2] = (0.5 * (a + b) + c) / h22
    hm[:, 1, 0] = d * i2h / h22
    hm[:, 1, 1] = e * i2h / h22
    hm[:, 1, 2] = (0.5 * (d + e) + f) / h22
    hm[:, 2, 0] = g * i2h / h22
    hm[:, 2, 1] = hh * i2h / h22
    hm[:, 2, 2] = 1.0
    return hm


def _rotations_from_h(hm):
    """IPPE core: the two candidate rotations from the homography's first-order
    expansion at the origin."""
    p, q = hm[:, 0, 2], hm[:, 1, 2]
    j00 = hm[:, 0, 0] - hm[:, 2, 0] * p
    j01 = hm[:, 0, 1] - hm[:, 2, 1] * p
    j10 = hm[:, 1, 0] - hm[:, 2, 0] * q
    j11 = hm[:, 1, 1] - hm[:, 2, 1] * q
    # Rv: minimal rotation taking the z axis onto (p, q, 1)/|.|
    s = np.sqrt(p * p + q * q + 1.0)
    ax, ay, az = p / s, q / s, 1.0 / s
    w = 1.0 / (1.0 + az)
    bsz = p.shape[0]
    rv = np.empty((bsz, 3, 3))
    rv[:, 0, 0] = 1.0 - ax * ax * w
    rv[:, 0, 1] = -ax * ay * w
    rv[:, 0, 2] = ax
    rv[:, 1, 0] = -ax * ay * w
    rv[:, 1, 1] = 1.0 - ay * ay * w
    rv[:, 1, 2] = ay
    rv[:, 2, 0] = -ax
    rv[:, 2, 1] = -ay
    rv[:, 2, 2] = az
    b00 = rv[:, 0, 0] - p * rv[:, 2, 0]
    b01 = rv[:, 0, 1] - p * rv[:, 2, 1]
    b10 = rv[:, 1, 0] - q * rv[:, 2, 0]
    b11 = rv[:, 1, 1] - q * rv[:, 2, 1]
    dtinv = 1.0 / (b00 * b11 - b01 * b10)
    bi00, bi01, bi10, bi11 = dtinv * b11, -dtinv * b01, -dtinv * b10, dtinv * b00
    a00 = bi00 * j00 + bi01 * j10
    a01 = bi00 * j01 + bi01 * j11
    a10 = bi10 * j00 + bi11 * j10
    a11 = bi10 * j01 + bi11 * j11
    ata00 = a00 * a00 + a01 * a01
    ata01 = a00 * a10 + a01 * a11
    ata11 = a10 * a10 + a11 * a11
    gamma = np.sqrt(0.5 * (ata00 + ata11 + np.sqrt((ata00 - ata11) ** 2 + 4.0 * ata01 * ata01)))
    r00, r01, r10, r11 = a00 / gamma, a01 / gamma, a10 / gamma, a11 / gamma
    b0 = np.sqrt(np.maximum(1.0 - r00 * r00 - r10 * r10, 0.0))
    b1 = np.sqrt(np.maximum(1.0 - r01 * r01 - r11 * r11, 0.0))
    sp = -r00 * r01 - r10 * r11
    b1 = np.where(sp < 0, -b1, b1)

    def full(sign):
        c0 = np.stack([r00, r10, sign * b0], axis=1)
        c1 = np.stack([r01, r11, sign * b1], axis=1)
        c2 = np.cross(c0, c1)
        rt = np.stack([c0, c1, c2], axis=2)
        return rv @ rt

    return full(1.0), full(-1.0)


def _translation(rm, x, y, half):
    """Least-squares t for fixed R from the 8 equations
    (R X_i + t)_xy - (u_i, v_i)(R X_i + t)_z = 0 (3x3 normal equations, adjugate)."""
    ox = np.array([-half, half, half, -half])
    oy = np.array([half, half, -half, -half])
    rx = rm[:, 0, 0:1] * ox + rm[:, 0, 1:2] * oy
    ry = rm[:, 1, 0:1] * ox + rm[:, 1, 1:2] * oy
    rz = rm[:, 2, 0:1] * ox + rm[:, 2, 1:2] * oy
    a2, b2 = -x, -y
    n = 4.0
    ata02 = a2.sum(1)
    ata12 = b2.sum(1)
    ata22 = (a2 * a2 + b2 * b2).sum(1)
    bx = -a2 * rz - rx
    by = -b2 * rz - ry
    atb0 = bx.sum(1)
    atb1 = by.sum(1)
    atb2 = (a2 * bx + b2 * by).sum(1)
    ata20, ata21 = ata02, ata12
    det_inv = 1.0 / (n * n * ata22 - n * ata12 * ata21 - ata02 * n * ata20)
    s00 = n * ata22 - ata12 * ata21
    s01 = ata02 * ata21
    s02 = -ata02 * n
    s10 = ata12 * ata20
    s11 = n * ata22 - ata02 * ata20
    s12 = -n * ata12
    s20 = -n * ata20
    s21 = -n * ata21
    s22 = n * n
    t0 = det_inv * (s00 * atb0 + s01 * atb1 + s02 * atb2)
    t1 = det_inv * (s10 * atb0 + s11 * atb1 + s12 * atb2)
    t2 = det_inv * (s20 * atb0 + s21 * atb1 + s22 * atb2)
    return np.stack([t0, t1, t2], axis=1)


def _reproj_error(rm, t, x, y, half):
    """Summed squared reprojection error in normalised coordinates; OpenCV holds
    the per-point differences and the sum as ``float``."""
    ox = np.array([-half, half, half, -half])
    oy = np.array([half, half, -half, -half])
    px = rm[:, 0, 0:1] * ox + rm[:, 0, 1:2] * oy + t[:, 0:1]
    py = rm[:, 1, 0:1] * ox + rm[:, 1, 1:2] * oy + t[:, 1:2]
    pz = rm[:, 2, 0:1] * ox + rm[:, 2, 1:2] * oy + t[:, 2:3]
    dx = (px / pz - x).astype(np.float32)
    dy = (py / pz - y).astype(np.float32)
    err = np.zeros(x.shape[0], dtype=np.float32)
    for i in range(4):
        err = err + (dx[:, i] * dx[:, i] + dy[:, i] * dy[:, i])
    return err


def rodrigues_vec(rm):
    """Rotation matrix -> rotation vector (``cv2.Rodrigues`` matrix branch)."""
    rx = rm[:, 2, 1] - rm[:, 1, 2]
    ry = rm[:, 0, 2] - rm[:, 2, 0]
    rz = rm[:, 1, 0] - rm[:, 0, 1]
    s = np.sqrt((rx * rx + ry * ry + rz * rz) * 0.25)
    c = np.clip((rm[:, 0, 0] + rm[:, 1, 1] + rm[:, 2, 2] - 1.0) * 0.5, -1.0, 1.0)
    theta = np.arccos(c)
    out = np.empty((rm.shape[0], 3))
    small = s < 1e-5
    with np.errstate(divide="ignore", invalid="ignore"):
        vth = theta / (2.0 * s)
    out[:, 0], out[:, 1], out[:, 2] = rx * vth, ry * vth, rz * vth
    if small.any():
        idx = np.flatnonzero(small)
        for i in idx:
            if c[i] > 0:
                out[i] = 0.0
            else:
                m = rm[i]
                t0 = np.sqrt(max((m[0, 0] + 1) * 0.5, 0.0))
                t1 = np.sqrt(max((m[1, 1] + 1) * 0.5, 0.0)) * (-1.0 if m[0, 1] < 0 else 1.0)
                t2 = np.sqrt(max((m[2, 2] + 1) * 0.5, 0.0)) * (-1.0 if m[0, 2] < 0 else 1.0)
                if abs(t0) < abs(t1) and abs(t0) < abs(t2) and ((m[1, 2] > 0) != (t1 * t2 > 0)):
                    t2 = -t2
                v = np.array([t0, t1, t2])
                out[i] = v * (theta[i] / np.sqrt(v @ v))
    return out


def rodrigues_matrix(rvec):
    """Rotation vector -> matrix (for comparing rotations instead of raw rvecs)."""
    rvec = np.atleast_2d(rvec)
    th = np.sqrt((rvec * rvec).sum(1))
    out = np.tile(np.eye(3), (rvec.shape[0], 1, 1))
    nz = th > 0
    k = np.zeros_like(rvec)
    k[nz] = rvec[nz] / th[nz, None]
    kx = np.zeros((rvec.shape[0], 3, 3))
    kx[:, 0, 1], kx[:, 0, 2] = -k[:, 2], k[:, 1]
    kx[:, 1, 0], kx[:, 1, 2] = k[:, 2], -k[:, 0]
    kx[:, 2, 0], kx[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    return out + s * kx + (1 - c) * (kx @ kx)


def ippe_square_both(corners, camera_matrix, dist_coeffs, marker_size):
    """Both IPPE solutions, ordered by reprojection error (smaller first).
    Returns dict(R1, t1, e1, R2, t2, e2)."""
    corners = np.asarray(corners, dtype=np.float32).reshape(-1, 4, 2)
    half = float(np.float32(marker_size / 2))  # object points are float32 (base_filter.py:112-120)
    x, y = undistort_normalize(corners, camera_matrix, dist_coeffs)
    hm = _homography_square(x, y, half)
    ra, rb = _rotations_from_h(hm)
    ta = _translation(ra, x, y, half)
    tb = _translation(rb, x, y, half)
    ea = _reproj_error(ra, ta, x, y, half)
    eb = _reproj_error(rb, tb, x, y, half)
    swap = ~(ea < eb)
    r1 = np.where(swap[:, None, None], rb, ra)
    r2 = np.where(swap[:, None, None], ra, rb)
    t1 = np.where(swap[:, None], tb, ta)
    t2 = np.where(swap[:, None], ta, tb)
    e1 = np.where(swap, eb, ea)
    e2 = np.where(swap, ea, eb)
    return dict(R1=r1, t1=t1, e1=e1, R2=r2, t2=t2, e2=e2)


def estimate_pose_of_markers(corners, camera_matrix, dist_coeffs, marker_size=0.16):
    """(B,4,2) float32 -> (B,6) float64 ``[tvec, rvec]`` (base_filter.py:92-171)."""
    sol = ippe_square_both(corners, camera_matrix, dist_coeffs, marker_size)
    return np.hstack([sol["t1"], rodrigues_vec(sol["R1"])])
