"""Sweep-kernel experiments on a 2000-marker EKF (N = 6010): time k_sweep alone."""
import ctypes as C, sys
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, _lib, synth
km, dist = synth.reference_calibration()
f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist, max_landmarks=2000)
f.set_landmarks(np.arange(2000))
lib = _lib.lib()
for cfg in (0, 1):
    for dbg in [int(a) for a in sys.argv[1:]] or [0]:
        us = C.c_double()
        _lib.check(lib.asb_bench_sweep(f._h, 60, cfg, dbg, 50, C.byref(us)))
        print(f"config {cfg} dbg {dbg:2d}: {us.value:8.2f} us", flush=True)
