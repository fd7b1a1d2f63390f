"""Serpentine tile walk + L2 eviction priorities for k_sweep (N = 6010, m4 = 60, dual config)."""
import ctypes as C, sys
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, _lib, synth
km, dist = synth.reference_calibration()
f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist, max_landmarks=2000)
f.set_landmarks(np.arange(2000))
lib = _lib.lib()
ALT = 1 << 24
def pol(ld, st, mir): return (ld | (st << 3) | (mir << 6)) << 12
cases = [("baseline .cg loads", 2 << 4),
         ("serpentine, .cg loads", ALT | (2 << 4)),
         ("serpentine, ld cg / st default / mirror evict-first", ALT | pol(2, 0, 4)),
         ("serpentine, ld cg / st evict-last / mirror evict-first", ALT | pol(2, 3, 4)),
         ("serpentine, ld evict-first / st evict-last / mirror evict-first", ALT | pol(4, 3, 4)),
         ("serpentine, ld cg / st evict-last / mirror .cs", ALT | pol(2, 3, 1)),
         ("forward, ld cg / st evict-last / mirror evict-first", pol(2, 3, 4)),
         ("serpentine, ld cg / st evict-last / mirror default", ALT | pol(2, 3, 0))]
for cfg in (1, 0):
    for name, dbg in cases:
        us = C.c_double()
        _lib.check(lib.asb_bench_sweep(f._h, 60, cfg, dbg, 50, C.byref(us)))
        print(f"config {cfg} {name:68s}: {us.value:8.2f} us", flush=True)
