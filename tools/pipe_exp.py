"""C3 stream (N = 6010, 20 markers/frame): frame time and the k_f1 / k_f2 timeline of the last frame."""
import ctypes as C, sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from aruco_slam_b200 import EKF, _lib, synth
km, dist = synth.reference_calibration()
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 100
stream, n_build = synth.make_map_build_stream(2000, 20, 20 + frames)
f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist, quat_mode="as_written", max_landmarks=2000)
ids, poses, off = stream.packed()
f.stream_upload(ids, poses, off)
f.stream_run(0, n_build + 20)
f.synchronize()
lib = _lib.lib()
t0 = time.perf_counter()
f.stream_run(n_build + 20, n_build + 20 + frames)
t1 = time.perf_counter()
f.synchronize()
t2 = time.perf_counter()
print(f"{frames} frames: enqueue {1e6 * (t1 - t0) / frames:.1f} us/frame, total {1e6 * (t2 - t0) / frames:.1f} us/frame")
us = (C.c_double * 12)()
_lib.check(lib.asb_front_profile(f._h, us))
print("k_f1 [T+issue, model+wait, lazy+Q+S, chol, outputs] us:", " ".join(f"{u:6.2f}" for u in us[:5]))
print(f"k_f1 total {us[5]:.2f}; k_f1 start -> k_f2 start {us[6]:.2f}, -> k_f2 end {us[7]:.2f}; k_f2 [loads, product, stores] {us[8]:.2f} {us[9]:.2f} {us[10]:.2f}")
tl = (C.c_double * 32)()
_lib.check(lib.asb_timeline(f._h, tl))
names = {0: "f1 start", 1: "f1 issued", 2: "f1 model", 3: "f1 S", 4: "f1 chol", 5: "f1 end", 6: "f1 lazy", 7: "f1 Q+I", 8: "f2 start", 9: "f2 loaded", 10: "f2 product", 11: "f2 end", 12: "sweep start", 13: "sweep cta0 end"}
for u in range(2):
    print(f"update {u}: " + ", ".join(f"{names[i]} {tl[16 * u + i]:.1f}" for i in sorted(names)))
import os
if os.environ.get("ASB_CTA_PROF"):
    n = 296
    cp = (C.c_double * (3 * n))()
    _lib.check(lib.asb_cta_profile(f._h, cp, n))
    a = np.array(cp[:]).reshape(n, 3)
    a = a[a[:, 0] >= 0]
    print(f"sweep CTAs {len(a)}: start min/max {a[:,0].min():.1f}/{a[:,0].max():.1f}, end min/median/max {a[:,1].min():.1f}/{np.median(a[:,1]):.1f}/{a[:,1].max():.1f}")
    dur = a[:, 1] - a[:, 0]
    print("duration percentiles 0/10/50/90/100:", np.percentile(dur, [0, 10, 50, 90, 100]).round(1))
    sm, cnt = np.unique(a[:, 2], return_counts=True)
    print("SMs used", len(sm), "CTAs per SM histogram", np.bincount(cnt))
    print("first 12 CTAs (start, end, sm):", a[:12].round(1).tolist())
    if len(a) == 296:
        smid = a[:, 2].astype(int)
        same = int((smid[:148] == smid[148:]).sum())
        print("CTA i and CTA 148+i on the same SM:", same, "of 148; sm[147], sm[295] =", smid[147], smid[295], "; sm[146], sm[294] =", smid[146], smid[294])
