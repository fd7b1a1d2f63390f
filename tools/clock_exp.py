"""SM clock / power while the sweep kernel runs back to back (is the combined MMA + HBM load power-capped?)."""
import ctypes as C, subprocess, sys, threading, time
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, _lib, synth
km, dist = synth.reference_calibration()
f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist, max_landmarks=2000)
f.set_landmarks(np.arange(2000))
lib = _lib.lib()
Q = "clocks.sm,clocks.mem,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown"
for cfg, dbg in ((1, 0), (1, 14), (1, 1), (2, 0), (2, 14), (2, 1)):
    rows = []
    p = subprocess.Popen(["nvidia-smi", f"--query-gpu={Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", "0"],
                         stdout=subprocess.PIPE, text=True)
    th = threading.Thread(target=lambda: [rows.append(l.strip()) for l in p.stdout], daemon=True)
    th.start()
    us = C.c_double()
    t0 = time.time()
    for _ in range(3):
        _lib.check(lib.asb_bench_sweep(f._h, 60, cfg, dbg, 8000, C.byref(us)))
    wall = time.time() - t0
    time.sleep(0.1)
    p.terminate()
    mid = rows[len(rows) // 3:]
    print(f"config {cfg} dbg {dbg:2d}: {us.value:7.2f} us/launch over {wall:.1f} s | last samples: {mid[-3:]}", flush=True)
