mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_single.py tests/test_gpu_batch_ippe.py -m gpu -q -x 2>&1 | tail -5
python - <<'PY'
import time, numpy as np
from aruco_slam_b200 import EKF, synth
km, dist = synth.reference_calibration()
f = EKF(np.array([0,0,0,1,0,0,0,0,0,0.0]), camera_matrix=km, dist_coeffs=dist, max_landmarks=64)
g = EKF(np.array([0,0,0,1,0,0,0,0,0,0.0]), camera_matrix=km, dist_coeffs=dist, max_landmarks=64)
c = synth.make_corner_sets(64, km, dist, seed=3)
ids = np.arange(20, dtype=np.int32)
for name, fn in (("fused process_detections", lambda: f.process_detections(c[:20], ids, True, 0, 0.16)),
                 ("separate estimate+observe+get_poses", lambda: (g.observe(ids, g.estimate_pose_of_markers(c[:20], ids, 0.16)), g.get_poses()))):
    for _ in range(20): fn()
    t0 = time.perf_counter()
    for _ in range(300): fn()
    print(f"{name}: {1e6*(time.perf_counter()-t0)/300:.1f} us per frame (20 markers, 20-landmark map)")
PY
