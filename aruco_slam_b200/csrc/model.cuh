// Measurement model and quaternion arithmetic of the two reference filters, as
// closed forms.  Device-only, FP64, shared
// by the single-filter engine (filter_single.cu) and the batched engine
// (filter_batch.cu).
//
// Reference (relative to the reference tree):
//   EKF  h / Jacobian   filters/extended_kalman_filter.py:302-350 (sympy lambdas)
//   ROT  h / Jacobian   filters/ekf_with_rotations.py:354-422
//   quaternion reset    extended_kalman_filter.py:138-152, ekf_with_rotations.py:146-177
//   add_marker          extended_kalman_filter.py:264-274, ekf_with_rotations.py:300-319
#pragma once
#include <cuda_runtime.h>

namespace asb {

constexpr int CAM = 10;  // x y z qw qx qy qz ex ey ez

struct Noise {
    double init_cam, init_lm, r, q_cam, q_err, q_lm;
};

template <int KIND> struct Dims;
template <> struct Dims<0> { static constexpr int D = 3, R = 3; };    // EKF
template <> struct Dims<1> { static constexpr int D = 10, R = 7; };   // EKF_Rotations

// p (x) q, scalar first.
__device__ __forceinline__ void hamilton(const double* p, const double* q, double* o) {
    o[0] = p[0] * q[0] - p[1] * q[1] - p[2] * q[2] - p[3] * q[3];
    o[1] = p[0] * q[1] + p[1] * q[0] + p[2] * q[3] - p[3] * q[2];
    o[2] = p[0] * q[2] - p[1] * q[3] + p[2] * q[0] + p[3] * q[1];
    o[3] = p[0] * q[3] + p[1] * q[2] - p[2] * q[1] + p[3] * q[0];
}

__device__ __forceinline__ void quat_normalize(double* q) {
    double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}

// L(p): p (x) q = L(p) q.   Rm(q): p (x) q = Rm(q) p.
__device__ __forceinline__ void left_matrix(const double* p, double m[4][4]) {
    m[0][0] = p[0]; m[0][1] = -p[1]; m[0][2] = -p[2]; m[0][3] = -p[3];
    m[1][0] = p[1]; m[1][1] = p[0];  m[1][2] = -p[3]; m[1][3] = p[2];
    m[2][0] = p[2]; m[2][1] = p[3];  m[2][2] = p[0];  m[2][3] = -p[1];
    m[3][0] = p[3]; m[3][1] = -p[2]; m[3][2] = p[1];  m[3][3] = p[0];
}
__device__ __forceinline__ void right_matrix(const double* q, double m[4][4]) {
    m[0][0] = q[0]; m[0][1] = -q[1]; m[0][2] = -q[2]; m[0][3] = -q[3];
    m[1][0] = q[1]; m[1][1] = q[0];  m[1][2] = q[3];  m[1][3] = -q[2];
    m[2][0] = q[2]; m[2][1] = -q[3]; m[2][2] = q[0];  m[2][3] = q[1];
    m[3][0] = q[3]; m[3][1] = q[2];  m[3][2] = -q[1]; m[3][3] = q[0];
}

// Homogeneous rotation quadratic form M(u) (sympy Quaternion.to_rotation_matrix
// before its 1/|u|^2 factor).
__device__ __forceinline__ void quad_form(const double* u, double m[3][3]) {
    const double a = u[0], b = u[1], c = u[2], d = u[3];
    m[0][0] = a * a + b * b - c * c - d * d; m[0][1] = 2 * (b * c - d * a); m[0][2] = 2 * (b * d + c * a);
    m[1][0] = 2 * (b * c + d * a); m[1][1] = a * a - b * b + c * c - d * d; m[1][2] = 2 * (c * d - b * a);
    m[2][0] = 2 * (b * d - c * a); m[2][1] = 2 * (c * d + b * a); m[2][2] = a * a - b * b - c * c + d * d;
}

// scipy Rotation.from_quat(q, scalar_first=True).as_matrix(): normalise, then M.
__device__ __forceinline__ void rotation_matrix(const double* q, double m[3][3]) {
    double u[4] = {q[0], q[1], q[2], q[3]};
    quat_normalize(u);
    quad_form(u, m);
}

// (Rotation.from_quat(dq) * Rotation.from_quat(q)).as_quat(), all scalar-first.
__device__ __forceinline__ void compose_normalized(const double* dq, const double* q, double* o) {
    double a[4] = {dq[0], dq[1], dq[2], dq[3]};
    double b[4] = {q[0], q[1], q[2], q[3]};
    quat_normalize(a);
    quat_normalize(b);
    hamilton(a, b, o);
    quat_normalize(o);
}

// Rotation.from_euler('xyz', a).as_quat(scalar_first=True): qz (x) qy (x) qx, no sign fix.
__device__ __forceinline__ void euler_xyz_quat(const double* ang, double* o) {
    double sx, cx, sy, cy, sz, cz;
    sincos(0.5 * ang[0], &sx, &cx);
    sincos(0.5 * ang[1], &sy, &cy);
    sincos(0.5 * ang[2], &sz, &cz);
    const double qx[4] = {cx, sx, 0.0, 0.0}, qy[4] = {cy, 0.0, sy, 0.0}, qz[4] = {cz, 0.0, 0.0, sz};
    double t[4];
    hamilton(qy, qx, t);
    hamilton(qz, t, o);
    quat_normalize(o);
}

// Rotation.from_matrix(m).as_quat(scalar_first=True): largest of (m00,m11,m22,trace),
// first maximum wins, normalised, sign not canonicalised.
__device__ __forceinline__ void matrix_to_quat(const double m[3][3], double* o) {
    const double tr = m[0][0] + m[1][1] + m[2][2];
    int choice = 0;
    double best = m[0][0];
    if (m[1][1] > best) { best = m[1][1]; choice = 1; }
    if (m[2][2] > best) { best = m[2][2]; choice = 2; }
    if (tr > best) { choice = 3; }
    double x, y, z, w;
    if (choice == 3) {
        x = m[2][1] - m[1][2]; y = m[0][2] - m[2][0]; z = m[1][0] - m[0][1]; w = 1 + tr;
    } else if (choice == 0) {
        x = 1 - tr + 2 * m[0][0]; y = m[1][0] + m[0][1]; z = m[2][0] + m[0][2]; w = m[2][1] - m[1][2];
    } else if (choice == 1) {
        x = m[1][0] + m[0][1]; y = 1 - tr + 2 * m[1][1]; z = m[2][1] + m[1][2]; w = m[0][2] - m[2][0];
    } else {
        x = m[2][0] + m[0][2]; y = m[2][1] + m[1][2]; z = 1 - tr + 2 * m[2][2]; w = m[1][0] - m[0][1];
    }
    o[0] = w; o[1] = x; o[2] = y; o[3] = z;
    quat_normalize(o);
}

// Position rows: h = M(u)^T (l - c)/|u|^2, u = (1,e) (x) q, and the 3x10 camera block /
// 3x3 landmark block of its Jacobian.  ju (3x4) = dh/du is returned for the ROT model.
__device__ __forceinline__ void position_model(const double* cam, const double* l, double h[3],
                                               double hc[3][CAM], double hl[3][3]) {
    const double p[4] = {1.0, cam[7], cam[8], cam[9]};
    const double* q = cam + 3;
    double u[4];
    hamilton(p, q, u);
    const double n = u[0] * u[0] + u[1] * u[1] + u[2] * u[2] + u[3] * u[3];
    const double inv_n = 1.0 / n;
    double m[3][3];
    quad_form(u, m);
    const double d[3] = {l[0] - cam[0], l[1] - cam[1], l[2] - cam[2]};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        h[r] = (m[0][r] * d[0] + m[1][r] * d[1] + m[2][r] * d[2]) * inv_n;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            hl[r][c] = m[c][r] * inv_n;
            hc[r][c] = -m[c][r] * inv_n;
        }
    }
    const double a = u[0], b = u[1], c = u[2], e = u[3];
    // (dM/du_k)^T d, k = 0..3
    double g[4][3];
    g[0][0] = 2 * (a * d[0] + e * d[1] - c * d[2]);
    g[0][1] = 2 * (-e * d[0] + a * d[1] + b * d[2]);
    g[0][2] = 2 * (c * d[0] - b * d[1] + a * d[2]);
    g[1][0] = 2 * (b * d[0] + c * d[1] + e * d[2]);
    g[1][1] = 2 * (c * d[0] - b * d[1] + a * d[2]);
    g[1][2] = 2 * (e * d[0] - a * d[1] - b * d[2]);
    g[2][0] = 2 * (-c * d[0] + b * d[1] - a * d[2]);
    g[2][1] = 2 * (b * d[0] + c * d[1] + e * d[2]);
    g[2][2] = 2 * (a * d[0] + e * d[1] - c * d[2]);
    g[3][0] = 2 * (-e * d[0] + a * d[1] + b * d[2]);
    g[3][1] = 2 * (-a * d[0] - e * d[1] + c * d[2]);
    g[3][2] = 2 * (b * d[0] + c * d[1] + e * d[2]);
    double ju[3][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int r = 0; r < 3; ++r) ju[r][k] = (g[k][r] - 2.0 * u[k] * h[r]) * inv_n;
    double lp[4][4], rq[4][4];
    left_matrix(p, lp);
    right_matrix(q, rq);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
            hc[r][3 + cc] = ju[r][0] * lp[0][cc] + ju[r][1] * lp[1][cc] + ju[r][2] * lp[2][cc] + ju[r][3] * lp[3][cc];
#pragma unroll
        for (int cc = 0; cc < 3; ++cc)
            hc[r][7 + cc] = ju[r][0] * rq[0][1 + cc] + ju[r][1] * rq[1][1 + cc] + ju[r][2] * rq[2][1 + cc] + ju[r][3] * rq[3][1 + cc];
    }
}

// Orientation rows of EKF_Rotations: q_cl = inv(u_c) (x) u_l; 4x7 camera block (state
// cols 3:10) and 4x7 landmark block (landmark cols 3:10).
__device__ __forceinline__ void orientation_model(const double* cam, const double* lm, double h[4],
                                                  double hc[4][7], double hl[4][7]) {
    const double pc[4] = {1.0, cam[7], cam[8], cam[9]};
    const double pl[4] = {1.0, lm[7], lm[8], lm[9]};
    const double* qc = cam + 3;
    const double* ql = lm + 3;
    double uc[4], ul[4];
    hamilton(pc, qc, uc);
    hamilton(pl, ql, ul);
    const double nc = uc[0] * uc[0] + uc[1] * uc[1] + uc[2] * uc[2] + uc[3] * uc[3];
    const double inc = 1.0 / nc;
    const double conj[4] = {uc[0], -uc[1], -uc[2], -uc[3]};
    const double inv_uc[4] = {conj[0] * inc, conj[1] * inc, conj[2] * inc, conj[3] * inc};
    hamilton(inv_uc, ul, h);
    double dul[4][4];  // d q_cl / d u_l = L(inv u_c)
    left_matrix(inv_uc, dul);
    double rul[4][4];  // Rm(u_l)
    right_matrix(ul, rul);
    // d inv(u)/du = (C - 2 conj u^T / n) / n, C = diag(1,-1,-1,-1)
    double dinv[4][4];
    const double cd[4] = {1.0, -1.0, -1.0, -1.0};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            dinv[i][j] = ((i == j ? cd[i] : 0.0) - 2.0 * conj[i] * uc[j] * inc) * inc;
    double duc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            duc[i][j] = rul[i][0] * dinv[0][j] + rul[i][1] * dinv[1][j] + rul[i][2] * dinv[2][j] + rul[i][3] * dinv[3][j];
    double lpc[4][4], rqc[4][4], lpl[4][4], rql[4][4];
    left_matrix(pc, lpc);
    right_matrix(qc, rqc);
    left_matrix(pl, lpl);
    right_matrix(ql, rql);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            hc[i][j] = duc[i][0] * lpc[0][j] + duc[i][1] * lpc[1][j] + duc[i][2] * lpc[2][j] + duc[i][3] * lpc[3][j];
            hl[i][j] = dul[i][0] * lpl[0][j] + dul[i][1] * lpl[1][j] + dul[i][2] * lpl[2][j] + dul[i][3] * lpl[3][j];
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            hc[i][4 + j] = duc[i][0] * rqc[0][1 + j] + duc[i][1] * rqc[1][1 + j] + duc[i][2] * rqc[2][1 + j] + duc[i][3] * rqc[3][1 + j];
            hl[i][4 + j] = dul[i][0] * rql[0][1 + j] + dul[i][1] * rql[1][1 + j] + dul[i][2] * rql[2][1 + j] + dul[i][3] * rql[3][1 + j];
        }
    }
}

// One observation: innovation nu = z - h (R rows) and the dense row block
// [R x 10 camera | R x D landmark], row-major with row length 10 + D.
template <int KIND>
__device__ __forceinline__ void observation_model(const double* cam, const double* lm,
                                                  const double* pose, double* nu, double* blk) {
    constexpr int D = Dims<KIND>::D, R = Dims<KIND>::R, W = CAM + D;
    double h3[3], hc[3][CAM], hl[3][3];
    position_model(cam, lm, h3, hc, hl);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        nu[r] = pose[r] - h3[r];
#pragma unroll
        for (int c = 0; c < CAM; ++c) blk[r * W + c] = hc[r][c];
#pragma unroll
        for (int c = 0; c < D; ++c) blk[r * W + CAM + c] = (c < 3) ? hl[r][c] : 0.0;
    }
    if constexpr (KIND == 1) {
        double h4[4], oc[4][7], ol[4][7], zq[4];
        orientation_model(cam, lm, h4, oc, ol);
        euler_xyz_quat(pose + 3, zq);  // z = [t, quat(from_euler('xyz', rvec))] (ROT:216-224)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            nu[3 + r] = zq[r] - h4[r];
            double* row = blk + (3 + r) * W;
            row[0] = row[1] = row[2] = 0.0;
#pragma unroll
            for (int c = 0; c < 7; ++c) row[3 + c] = oc[r][c];
            row[CAM + 0] = row[CAM + 1] = row[CAM + 2] = 0.0;
#pragma unroll
            for (int c = 0; c < 7; ++c) row[CAM + 3 + c] = ol[r][c];
        }
    }
}

// add_marker (EKF:264-274, ROT:300-319): new landmark from the camera state and an
// observation row.  NOTE the reference places it with R(q)^-1, the same rotation h uses
//  — replicated.
template <int KIND>
__device__ __forceinline__ void new_landmark(const double* cam, const double* pose, double* lm) {
    double rm[3][3];
    rotation_matrix(cam + 3, rm);  // rot_mc; rot_cm = rm^T
#pragma unroll
    for (int r = 0; r < 3; ++r)
        lm[r] = rm[0][r] * pose[0] + rm[1][r] * pose[1] + rm[2][r] * pose[2] + cam[r];
    if constexpr (KIND == 1) {
        double ql[4], rcl[3][3], prod[3][3];
        euler_xyz_quat(pose + 3, ql);
        quad_form(ql, rcl);
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                prod[i][j] = rm[0][i] * rcl[0][j] + rm[1][i] * rcl[1][j] + rm[2][i] * rcl[2][j];
        matrix_to_quat(prod, lm + 3);
        lm[7] = lm[8] = lm[9] = 0.0;
    }
}

// Camera part of the state update (EKF:134-152 / ROT:143-157).
__device__ __forceinline__ void camera_reset(double* cam, const double* dx, int as_written) {
    cam[0] += dx[0]; cam[1] += dx[1]; cam[2] += dx[2];
    const double q[4] = {cam[3], cam[4], cam[5], cam[6]};
    const double dq[4] = {1.0, 0.5 * dx[7], 0.5 * dx[8], 0.5 * dx[9]};
    double o[4];
    if (as_written) {
        // from_quat() reads the scalar-first arrays as (x,y,z,w) (EKF:145-146):
        // P([a,b,c,d]) = (w=d, x=a, y=b, z=c); product stored back as [w,x,y,z].
        const double qp[4] = {q[3], q[0], q[1], q[2]};
        const double dqp[4] = {dq[3], dq[0], dq[1], dq[2]};
        compose_normalized(dqp, qp, o);
    } else {
        compose_normalized(dq, q, o);
    }
    cam[3] = o[0]; cam[4] = o[1]; cam[5] = o[2]; cam[6] = o[3];
    cam[7] = cam[8] = cam[9] = 0.0;
}

// Landmark part: EKF adds dx (EKF:135); ROT adds xyz and left-multiplies the quaternion
// by normalize([1, dx_err/2]); landmark error dims are never written (ROT:160-177).
template <int KIND>
__device__ __forceinline__ void landmark_update(double* lm, const double* dx) {
    lm[0] += dx[0]; lm[1] += dx[1]; lm[2] += dx[2];
    if constexpr (KIND == 1) {
        const double dq[4] = {1.0, 0.5 * dx[7], 0.5 * dx[8], 0.5 * dx[9]};
        double o[4];
        compose_normalized(dq, lm + 3, o);
        lm[3] = o[0]; lm[4] = o[1]; lm[5] = o[2]; lm[6] = o[3];
    }
}

__device__ __forceinline__ double process_noise(const Noise& nz, int i) {
    return i < 3 ? nz.q_cam : (i < 7 ? 0.0 : (i < CAM ? nz.q_err : nz.q_lm));
}

}  // namespace asb
