import sys, time
sys.path.insert(0, ".")
from aruco_slam_b200 import FilterBatch, synth
ids, poses, off = synth.make_batch(4096, 32, 300, 0.5, "ekf", seq_base=0, unique=64)
b = FilterBatch("ekf", 4096, 32, 32, quat_mode="as_written")
b.upload(ids, poses, off)
r = []
for rep in range(4):
    b.reset(); b.synchronize()
    t0 = time.perf_counter(); b.run(0, 300); b.synchronize(); r.append(int(off[-1]) / (time.perf_counter() - t0) / 1e6)
print("runs (M updates/s):", " ".join(f"{x:.1f}" for x in r))
