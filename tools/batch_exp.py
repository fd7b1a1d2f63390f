"""Phase profile of k_batch (CTA 0) on a reduced C2."""
import ctypes as C, sys, time
import numpy as np
sys.path.insert(0, ".")
from aruco_slam_b200 import FilterBatch, _lib, synth
nseq, frames = int(sys.argv[1]) if len(sys.argv) > 1 else 1184, int(sys.argv[2]) if len(sys.argv) > 2 else 100
ids, poses, off = synth.make_batch(nseq, 32, frames, 0.5, "ekf", seq_base=0, unique=64)
b = FilterBatch("ekf", nseq, 32, 32, quat_mode="as_written")
b.upload(ids, poses, off)
dts = []
for rep in range(int(sys.argv[3]) if len(sys.argv) > 3 else 2):
    b.reset(); b.synchronize()
    clk = (C.c_double * 10)()
    _lib.check(_lib.lib().asb_batch_profile(b._h, clk))
    t0 = time.perf_counter(); b.run(0, frames); b.synchronize(); dt = time.perf_counter() - t0
    dts.append(dt)
    _lib.check(_lib.lib().asb_batch_profile(b._h, clk))
print("runs (M updates/s):", " ".join(f"{int(off[-1])/d/1e6:.1f}" for d in dts))
dt = min(dts)
names = ["assoc", "augment+predict", "model", "G", "S", "chol", "solve", "-", "Pupdate", "apply"]
tot = sum(clk)
print(f"{nseq} seq x {frames} frames: {dt*1e3:.1f} ms, {int(off[-1])/dt/1e6:.1f} M updates/s; CTA0 clocks per frame:")
nfr = frames * ((nseq + 0) // max(1, min(nseq, 148 * 3)))
for n, c in zip(names, clk):
    print(f"  {n:16s} {c/ (frames):10.0f}  {100*c/tot:5.1f}%")
