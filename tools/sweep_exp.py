"""Sweep-kernel experiments on a 2000-marker EKF (N = 6010): time k_sweep alone."""
import ctypes as C, sys
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, _lib, synth
km, dist = synth.reference_calibration()
f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist, max_landmarks=2000)
f.set_landmarks(np.arange(2000))
lib = _lib.lib()
for cfg in (0, 1):   # 0 = 128 x 64 tiles, 1 CTA / SM; 1 = 64 x 64 tiles, 2 CTAs / SM (the product path)
    for dbg in [int(a) for a in sys.argv[1:]] or [0]:
        us = C.c_double()
        _lib.check(lib.asb_bench_sweep(f._h, 60, cfg, dbg, 50, C.byref(us)))
        print(f"config {cfg} dbg {dbg:2d}: {us.value:8.2f} us", flush=True)
# phases of k_f1 / k_f2 on a real frame
rng = np.random.default_rng(0)
for rep in range(3):
    ids = rng.choice(2000, 20, replace=False)
    poses = np.concatenate([rng.normal(0, 1, (20, 3)) + [0, 0, 5], rng.normal(0, 0.1, (20, 3))], 1)
    f.observe(ids, poses)
    us = (C.c_double * 12)()
    _lib.check(lib.asb_front_profile(f._h, us))
    print("k_f1 phases [index, copies, model, lazy+S, chol] us, k_f1 total, f1->f2 start, f1->f2 end | k_f2 [loads, product, stores]:", " ".join(f"{u:6.2f}" for u in us[:7]), "| chol step clks [diag panel bar trail bar]:", " ".join(f"{u:5.0f}" for u in us[7:12]), flush=True)
