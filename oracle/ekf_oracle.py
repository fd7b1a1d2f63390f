"""CPU oracle: numpy restatement of the reference's two multiplicative EKFs.

TEST INFRASTRUCTURE — not the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product path
(``aruco_slam_b200``) never does and fails loudly without its CUDA library.

What is restated (file:line relative to /root/reference):
  * ``EKF``            filters/extended_kalman_filter.py:37-357
  * ``EKF_Rotations``  filters/ekf_with_rotations.py:43-431
The sympy-generated ``h`` / Jacobian lambdas (extended_kalman_filter.py:292-353,
ekf_with_rotations.py:337-427) are replaced by their closed forms (SURVEY.md §8a
rows A3/A3r); scipy ``Rotation`` calls by ``oracle/rotations.py``; SuperLU
``spsolve`` (extended_kalman_filter.py:129) by dense LU ``numpy.linalg.solve``.

PARITY PINNING: the reference has no tests or golden vectors of its own
(SURVEY.md §4), so this oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF,
run in the build container by ``oracle/make_golden.py`` and committed under
``tests/golden/`` (checked by ``tests/test_oracle_golden.py``).
"""
from __future__ import annotations

import numpy as np

from . import rotations as rot

CAM_DIMS = 10

# extended_kalman_filter.py:21-27
EKF_NOISE = dict(init_cam=0.1, init_lm=0.7, r=0.9, q_cam=0.3, q_err=0.5, q_lm=0.01)
# ekf_with_rotations.py:23-31 (Q_UNCERTAINTY_LM_QUAT is never read)
ROT_NOISE = dict(init_cam=0.1, init_lm=0.7, r=0.9, q_cam=0.2, q_err=0.5, q_lm=0.01)


# ----------------------------------------------------------------------------
# measurement model, closed form
# ----------------------------------------------------------------------------
def h_jac_position(c, q, e, l):
    """h = R(u)^-1 (l - c), u = (1, e) ⊗ q, and its 3x13 Jacobian w.r.t.
    [c(3), q(4), e(3), l(3)]  (extended_kalman_filter.py:310-350)."""
    p = np.array([1.0, e[0], e[1], e[2]])
    u = rot.hamilton(p, q)
    n = u @ u
    m = rot.quad_form_matrix(u)
    d = np.asarray(l, dtype=np.float64) - np.asarray(c, dtype=np.float64)
    h = m.T @ d / n
    dm = rot.quad_form_derivs(u)
    j_u = np.empty((3, 4))
    for k in range(4):
        j_u[:, k] = (dm[k].T @ d - 2.0 * u[k] * h) / n
    jac = np.empty((3, 13))
    jac[:, 0:3] = -m.T / n
    jac[:, 3:7] = j_u @ rot.left_matrix(p)
    jac[:, 7:10] = j_u @ rot.right_matrix(q)[:, 1:4]
    jac[:, 10:13] = m.T / n
    return h, jac


def h_jac_pose(c, qc, ec, l, ql, el):
    """7-row model of EKF_Rotations (ekf_with_rotations.py:362-422):
    rows 0:3 as :func:`h_jac_position`; rows 3:7 q_cl = inv(u_c) ⊗ u_l.
    Jacobian 7x20 w.r.t. [c, qc, ec, l, ql, el]."""
    h3, j3 = h_jac_position(c, qc, ec, l)
    pc = np.array([1.0, ec[0], ec[1], ec[2]])
    pl = np.array([1.0, el[0], el[1], el[2]])
    uc = rot.hamilton(pc, qc)
    ul = rot.hamilton(pl, ql)
    nc = uc @ uc
    conj = np.array([uc[0], -uc[1], -uc[2], -uc[3]])
    inv_uc = conj / nc
    q_cl = rot.hamilton(inv_uc, ul)
    h = np.concatenate([h3, q_cl])

    jac = np.zeros((7, 20))
    jac[0:3, 0:10] = j3[:, 0:10]
    jac[0:3, 10:13] = j3[:, 10:13]
    # d q_cl / d u_l = L(inv u_c);   d q_cl / d u_c = Rm(u_l) . d inv(u_c)/d u_c
    d_ul = rot.left_matrix(inv_uc)
    cdiag = np.diag([1.0, -1.0, -1.0, -1.0])
    d_inv = (cdiag - 2.0 * np.outer(conj, uc) / nc) / nc
    d_uc = rot.right_matrix(ul) @ d_inv
    jac[3:7, 3:7] = d_uc @ rot.left_matrix(pc)
    jac[3:7, 7:10] = d_uc @ rot.right_matrix(qc)[:, 1:4]
    jac[3:7, 13:17] = d_ul @ rot.left_matrix(pl)
    jac[3:7, 17:20] = d_ul @ rot.right_matrix(ql)[:, 1:4]
    return h, jac


# ----------------------------------------------------------------------------
# filters
# ----------------------------------------------------------------------------
class _OracleBase:
    LM_DIMS = 3
    ROWS = 3
    NOISE = EKF_NOISE

    def __init__(self, initial_camera_pose=None, faithful_cov: bool = True):
        pose = [0, 0, 0, 1, 0, 0, 0, 0, 0, 0] if initial_camera_pose is None \
            else initial_camera_pose
        self.state = np.array(pose, dtype=np.float64)
        self.uncertainty = np.eye(CAM_DIMS) * self.NOISE["init_cam"]
        self.num_landmarks = 0
        self.landmarks = {}
        # faithful_cov=True: P <- (I - K H) @ P as a dense N x N product, exactly as
        # written (extended_kalman_filter.py:155-156).  False: P - K (H P).
        self.faithful_cov = faithful_cov
        self.last = {}

    # extended_kalman_filter.py:58-82 / ekf_with_rotations.py:66-90
    def observe(self, ids, poses):
        for idx, pose in zip(ids, poses):
            if idx in self.landmarks:
                continue
            self.add_marker(idx, pose)
        self.predict()
        self.update(ids, poses)

    def get_poses(self):
        return self.state[:CAM_DIMS], self.state[CAM_DIMS:].reshape(-1, self.LM_DIMS)

    def get_lm_uncertainties(self):
        return np.diagonal(self.uncertainty)[CAM_DIMS:].reshape(-1, self.LM_DIMS)

    def get_lm_estimates(self):
        return self.landmarks.items()

    # extended_kalman_filter.py:95-105 / ekf_with_rotations.py:103-113
    def predict(self):
        n = self.state.shape[0]
        q = np.zeros(n)
        q[0:3] = self.NOISE["q_cam"]
        q[7:10] = self.NOISE["q_err"]
        q[CAM_DIMS:] = self.NOISE["q_lm"]
        self.uncertainty = np.asarray(self.uncertainty) + np.diag(q)

    def _grow_cov(self, uncertainity):
        d = self.LM_DIMS
        n_dims = d * self.num_landmarks + CAM_DIMS
        new = np.eye(n_dims)
        if uncertainity is not None:
            new[-d:, -d:] *= uncertainity
        else:
            new[-d:, -d:] *= self.NOISE["init_lm"]
        new[: n_dims - d, : n_dims - d] = self.uncertainty
        self.uncertainty = new

    def _gain(self, z, h, dh):
        """extended_kalman_filter.py:120-131."""
        p = np.asarray(self.uncertainty)
        m = dh.shape[0]
        s = dh @ p @ dh.T + np.eye(m) * self.NOISE["r"]
        s_inv = np.linalg.solve(s, np.eye(m))
        k = p @ dh.T @ s_inv
        dx = k @ (z - h)
        self.last = dict(z=z, h=h, H=dh, S=s, K=k, dx=dx)
        return k, dx

    def _cov_update(self, k, dh):
        p = np.asarray(self.uncertainty)
        if self.faithful_cov:
            ident = np.eye(p.shape[0])
            self.uncertainty = (ident - k @ dh) @ p
        else:
            self.uncertainty = p - k @ (dh @ p)


class EKFOracle(_OracleBase):
    """``EKF`` (filters/extended_kalman_filter.py).  quat_mode:
    'as_written'   — lines 145-146 parse the scalar-first arrays as scalar-last
                     (SURVEY.md F2);
    'scalar_first' — those two calls with ``scalar_first=True``."""
    LM_DIMS = 3
    ROWS = 3
    NOISE = EKF_NOISE

    def __init__(self, initial_camera_pose=None, quat_mode="as_written", faithful_cov=True):
        super().__init__(initial_camera_pose, faithful_cov)
        if quat_mode not in ("as_written", "scalar_first"):
            raise ValueError(quat_mode)
        self.quat_mode = quat_mode

    # extended_kalman_filter.py:239-290
    def add_marker(self, idx, pose, uncertainity=None):
        self.landmarks[idx] = self.num_landmarks
        self.num_landmarks += 1
        rot_mc = rot.rotation_matrix(self.state[3:7])
        rot_cm = np.linalg.inv(rot_mc)
        t_ml = rot_cm @ np.asarray(pose[0:3], dtype=np.float64) + self.state[0:3]
        self.state = np.hstack((self.state, t_ml))
        self._grow_cov(uncertainity)

    # extended_kalman_filter.py:158-237
    def parse_poses(self, ids, poses):
        n = self.state.shape[0]
        k = len(ids)
        z = np.empty(3 * k)
        h = np.empty(3 * k)
        dh = np.zeros((3 * k, n))
        for j, (idx, pose) in enumerate(zip(ids, poses)):
            index = self.landmarks[idx]  # KeyError for unknown ids, as the reference
            s = 3 * index + CAM_DIMS
            hj, jac = h_jac_position(self.state[0:3], self.state[3:7], self.state[7:10],
                                     self.state[s:s + 3])
            z[3 * j:3 * j + 3] = pose[0:3]
            h[3 * j:3 * j + 3] = hj
            dh[3 * j:3 * j + 3, 0:CAM_DIMS] = jac[:, 0:CAM_DIMS]
            dh[3 * j:3 * j + 3, s:s + 3] = jac[:, CAM_DIMS:]
        return z, h, dh

    # extended_kalman_filter.py:107-156
    def update(self, ids, poses):
        z, h, dh = self.parse_poses(ids, poses)
        k, dx = self._gain(z, h, dh)
        self.state[0:3] += dx[0:3]
        self.state[CAM_DIMS:] += dx[CAM_DIMS:]
        q = self.state[3:7].copy()
        dq = np.array([1.0, *(dx[7:10] / 2)])
        if self.quat_mode == "as_written":
            # from_quat() reads [a,b,c,d] as (x=a, y=b, z=c, w=d); product; written
            # back scalar-first (extended_kalman_filter.py:145-149)
            qp = np.array([q[3], q[0], q[1], q[2]])
            dqp = np.array([dq[3], dq[0], dq[1], dq[2]])
            self.state[3:7] = rot.compose_normalized(dqp, qp)
        else:
            self.state[3:7] = rot.compose_normalized(dq, q)
        self.state[7:10] = 0.0
        self._cov_update(k, dh)


class EKFRotOracle(_OracleBase):
    """``EKF_Rotations`` (filters/ekf_with_rotations.py)."""
    LM_DIMS = 10
    ROWS = 7
    NOISE = ROT_NOISE

    # ekf_with_rotations.py:275-335
    def add_marker(self, idx, pose, uncertainity=None):
        self.landmarks[idx] = self.num_landmarks
        self.num_landmarks += 1
        rot_mc = rot.rotation_matrix(self.state[3:7])
        rot_cm = np.linalg.inv(rot_mc)
        # pose[QUAT_DIMS] = pose[3:7] of a 6-vector = the 3 rvec numbers, read as Euler xyz
        rot_cl = rot.euler_xyz_matrix(pose[3:6])
        t_ml = rot_cm @ np.asarray(pose[0:3], dtype=np.float64) + self.state[0:3]
        q_ml = rot.matrix_to_quat(rot_cm @ rot_cl)
        self.state = np.hstack((self.state, t_ml, q_ml, np.zeros(3)))
        self._grow_cov(uncertainity)

    # ekf_with_rotations.py:183-273
    def parse_poses(self, ids, poses):
        n = self.state.shape[0]
        k = len(ids)
        z = np.empty(7 * k)
        h = np.empty(7 * k)
        dh = np.zeros((7 * k, n))
        for j, (idx, pose) in enumerate(zip(ids, poses)):
            index = self.landmarks[idx]
            s = 10 * index + CAM_DIMS
            lm = self.state[s:s + 10]
            hj, jac = h_jac_pose(self.state[0:3], self.state[3:7], self.state[7:10],
                                 lm[0:3], lm[3:7], lm[7:10])
            z[7 * j:7 * j + 3] = pose[0:3]
            z[7 * j + 3:7 * j + 7] = rot.euler_xyz_quat(pose[3:6])
            h[7 * j:7 * j + 7] = hj
            dh[7 * j:7 * j + 7, 0:CAM_DIMS] = jac[:, 0:CAM_DIMS]
            dh[7 * j:7 * j + 7, s:s + 10] = jac[:, CAM_DIMS:]
        return z, h, dh

    # ekf_with_rotations.py:115-181
    def update(self, ids, poses):
        z, h, dh = self.parse_poses(ids, poses)
        k, dx = self._gain(z, h, dh)
        self.state[0:3] += dx[0:3]
        dq = np.array([1.0, *(dx[7:10] / 2)])
        self.state[3:7] = rot.compose_normalized(dq, self.state[3:7])
        self.state[7:10] = 0.0
        for i in range(self.num_landmarks):
            s = 10 * i + CAM_DIMS
            self.state[s:s + 3] += dx[s:s + 3]
            ldq = np.array([1.0, *(dx[s + 7:s + 10] / 2)])
            self.state[s + 3:s + 7] = rot.compose_normalized(ldq, self.state[s + 3:s + 7])
        self._cov_update(k, dh)


def make(kind: str, quat_mode: str = "as_written", initial_camera_pose=None, faithful_cov=True):
    if kind == "ekf":
        return EKFOracle(initial_camera_pose, quat_mode, faithful_cov)
    if kind in ("rot", "ekf_rotations"):
        return EKFRotOracle(initial_camera_pose, faithful_cov)
    raise ValueError(kind)
