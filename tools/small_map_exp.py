"""Per-frame observe() + get_poses() latency of the one-launch engine against the pipelined engine as the map grows
(chooses the default of asb_set_small_map_limit)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, EKF_Rotations, synth
km, dist = synth.reference_calibration()
pose0 = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
for cls, kind, sizes in ((EKF, "ekf", (10, 20, 30, 40, 50, 60)), (EKF_Rotations, "rot", (6, 10, 14, 18))):
    for nm in sizes:
        p = min(0.9, 10.0 / nm)
        st = synth.make_stream(77, nm, 400, p, kind)
        res = []
        for limit in (4096, 0):
            f = cls(pose0, camera_matrix=km, dist_coeffs=dist, max_landmarks=64)
            f.set_small_map_limit(limit)
            for t in range(100):
                f.observe(st.ids[t], st.poses[t]); f.get_poses()
            t0 = time.perf_counter()
            for t in range(100, 400):
                f.observe(st.ids[t], st.poses[t]); f.get_poses()
            res.append((time.perf_counter() - t0) / 300 * 1e6)
        d = 3 if kind == "ekf" else 10
        print(f"{kind} {nm} markers (N = {10 + d * nm}), k ~ {p * nm:.0f}: one launch {res[0]:.1f} us/frame, pipelined {res[1]:.1f} us/frame")
