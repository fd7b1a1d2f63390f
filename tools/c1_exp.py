"""Where a per-frame observe() + get_poses() goes at the C1 shape (1 sequence x 10 markers): wrapper vs library call."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, EKF_Rotations, _lib, synth
km, dist = synth.reference_calibration()
pose0 = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
for cls, kind in ((EKF, "ekf"), (EKF_Rotations, "rot")):
    st = synth.make_stream(1234, 10, 1000, 0.8, kind)
    f = cls(pose0, camera_matrix=km, dist_coeffs=dist)
    for t in range(50):
        f.observe(st.ids[t], st.poses[t]); f.get_poses()
    t_obs = t_get = 0.0
    for t in range(50, 1000):
        a = time.perf_counter(); f.observe(st.ids[t], st.poses[t]); b = time.perf_counter(); f.get_poses(); c = time.perf_counter()
        t_obs += b - a; t_get += c - b
    n = 950
    print(f"{kind}: observe {t_obs/n*1e6:.1f} us, get_poses {t_get/n*1e6:.1f} us per frame")
    # the library calls alone (same handle, arrays prepared once per frame outside the timed part)
    lib = _lib.lib(); h = f._h
    state = np.empty(10 + f.LM_DIMS * 64); nn = _lib.C.c_int(0) if hasattr(_lib, "C") else None
    t_lib = 0.0
    f2 = cls(pose0, camera_matrix=km, dist_coeffs=dist)
    for t in range(1000):
        ids = np.ascontiguousarray(st.ids[t], dtype=np.int32); poses = np.ascontiguousarray(st.poses[t], dtype=np.float64)
        a = time.perf_counter()
        lib.asb_observe(f2._h, _lib.iptr(ids), _lib.dptr(poses), ids.size)
        b = time.perf_counter()
        if t >= 50: t_lib += b - a
    print(f"{kind}: asb_observe alone (asynchronous return) {t_lib/n*1e6:.1f} us per frame")
    f2.get_poses()

# k_f1 / k_f2 phase times of the last update (globaltimer stamps, asb_front_profile)
import ctypes as C
for cls, kind in ((EKF, "ekf"), (EKF_Rotations, "rot")):
    st = synth.make_stream(1234, 10, 300, 0.8, kind)
    f = cls(pose0, camera_matrix=km, dist_coeffs=dist)
    acc = np.zeros(12); cnt = 0
    for t in range(300):
        f.observe(st.ids[t], st.poses[t])
        if t >= 50 and len(st.ids[t]) <= 8:
            us = (C.c_double * 12)()
            _lib.check(_lib.lib().asb_front_profile(f._h, us)); acc += np.array(us[:]); cnt += 1
    names = ["f1 index set + copy issue", "f1 model + copy wait", "f1 lazy operand + Q + S", "f1 cholesky", "f1 outputs (A = L^-1 H)", "f1 total", "f1 start->f2 start", "f1 start->f2 end", "f2 loads", "f2 product", "f2 stores", "-"]
    print(kind, "single-chunk frames:", ", ".join(f"{n} {v:.1f}" for n, v in zip(names, acc / max(cnt, 1))))
