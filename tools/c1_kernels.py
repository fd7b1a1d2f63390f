"""Short per-frame run at the C1 shape for an ncu launch list (kernel durations of the one-launch path)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, EKF_Rotations, synth
km, dist = synth.reference_calibration()
pose0 = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
for cls, kind in ((EKF, "ekf"), (EKF_Rotations, "rot")):
    st = synth.make_stream(1234, 10, 40, 0.8, kind)
    f = cls(pose0, camera_matrix=km, dist_coeffs=dist)
    for t in range(40):
        f.observe(st.ids[t], st.poses[t]); f.get_poses()
