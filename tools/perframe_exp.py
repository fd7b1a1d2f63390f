"""C3 (N = 6010, 20 markers / frame) through per-frame observe() calls: frame time with / without the pose
read-back, and the k_f1 / k_f2 / k_sweep timeline of the last two updates (asb_timeline)."""
import ctypes as C, sys, time
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import EKF, _lib, synth
km, dist = synth.reference_calibration()
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 60
stream, n_build = synth.make_map_build_stream(2000, 20, 20 + 4 * frames)
f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist, quat_mode="as_written", max_landmarks=2000)
ids, poses, off = stream.packed()
f.stream_upload(ids[:off[n_build + 20]], poses[:off[n_build + 20]], off[:n_build + 21])
f.stream_run(0, n_build + 20)
f.synchronize()
lib = _lib.lib()
t = n_build + 20
for mode in ("observe only", "observe + get_poses", "observe + synchronize"):
    f.synchronize()
    t0 = time.perf_counter()
    for _ in range(frames):
        f.observe(stream.ids[t], stream.poses[t])
        if mode == "observe + get_poses":
            cam, _lm = f.get_poses()
        elif mode == "observe + synchronize":
            f.synchronize()
        t += 1
    t1 = time.perf_counter()
    f.synchronize()
    t2 = time.perf_counter()
    print(f"{mode:24s}: enqueue {1e6 * (t1 - t0) / frames:7.1f} us/frame, total {1e6 * (t2 - t0) / frames:7.1f} us/frame")
    tl = (C.c_double * 32)()
    _lib.check(lib.asb_timeline(f._h, tl))
    names = {0: "f1 start", 5: "f1 end", 8: "f2 start", 11: "f2 end", 12: "sweep start", 13: "sweep cta0 end"}
    for u in range(2):
        print(f"   update {u}: " + ", ".join(f"{names[i]} {tl[16 * u + i]:.1f}" for i in sorted(names)))
# where does the host time go: the raw C call vs the Python wrapper
ids_l = [_lib.as_i32(stream.ids[t + i]) for i in range(frames)]
poses_l = [_lib.as_f64(stream.poses[t + i], (-1, 6)) for i in range(frames)]
f.synchronize()
call = []
t0 = time.perf_counter()
for i in range(frames):
    a = time.perf_counter()
    rc = lib.asb_observe(f._h, _lib.iptr(ids_l[i]), _lib.dptr(poses_l[i]), ids_l[i].size)
    call.append(time.perf_counter() - a)
t1 = time.perf_counter()
f.synchronize()
t2 = time.perf_counter()
call = np.array(call) * 1e6
print(f"raw asb_observe: per call min {call.min():.1f} median {np.median(call):.1f} max {call.max():.1f} us; loop {1e6 * (t1 - t0) / frames:.1f} us/frame, total {1e6 * (t2 - t0) / frames:.1f}")
print("first 12 calls (us):", call[:12].round(1).tolist())
