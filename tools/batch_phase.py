"""Phase profile of k_batch (CTA 0) for one shape: python tools/batch_phase.py kind n_seq n_markers frames p"""
import ctypes as C, sys, time
sys.path.insert(0, ".")
from aruco_slam_b200 import FilterBatch, _lib, synth
kind, nseq, nm, frames, p = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
ids, poses, off = synth.make_batch(nseq, nm, frames, p, kind, unique=min(nseq, 64))
b = FilterBatch(kind, nseq, nm, nm, quat_mode="scalar_first", engine="resident")
b.upload(ids, poses, off)
clk = (C.c_double * 10)()
for rep in range(2):
    b.reset(); b.synchronize()
    _lib.check(_lib.lib().asb_batch_profile(b._h, clk))
    t0 = time.perf_counter(); b.run(0, frames); b.synchronize(); dt = time.perf_counter() - t0
    _lib.check(_lib.lib().asb_batch_profile(b._h, clk))
names = ["assoc", "augment+predict", "model", "G", "S", "chol", "solve", "-", "Pupdate", "apply"]
tot = sum(clk)
print(f"{kind} {nseq} seq x {nm} markers x {frames} frames: {dt/frames*1e6:.1f} us/frame; CTA0 clocks per frame:")
for n, c in zip(names, clk):
    print(f"  {n:16s} {c/frames:10.0f}  {100*c/tot:5.1f}%")
