"""C4 through host buffers: repeated calls of the public IPPE entry point with page-locked buffers, next to the raw
PCIe ceiling (32 MB in + 48 MB out as two concurrent torch copies)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from aruco_slam_b200 import synth
from aruco_slam_b200.ippe import estimate_pose_of_markers
n = 1 << 20
km, dist = synth.reference_calibration()
corners = synth.make_corners(n, km, dist, 0.16, seed=3) if hasattr(synth, "make_corners") else None
if corners is None:
    rng = np.random.default_rng(3)
    c0 = rng.uniform(200, 800, size=(n, 1, 2)).astype(np.float32)
    sq = np.array([[0, 0], [60, 3], [58, 62], [-2, 57]], np.float32)
    corners = (c0 + sq[None] + rng.normal(0, 0.5, size=(n, 4, 2)).astype(np.float32)).astype(np.float32)
tc = torch.empty((n, 4, 2), dtype=torch.float32).pin_memory(); tc.numpy()[:] = corners.reshape(n, 4, 2)
to = torch.empty((n, 6), dtype=torch.float64).pin_memory()
hc, ho = tc.numpy(), to.numpy()
ts = []
for r in range(8):
    t0 = time.perf_counter(); estimate_pose_of_markers(hc, km, dist, 0.16, out=ho); ts.append(time.perf_counter() - t0)
print("e2e ms per call:", " ".join(f"{t*1e3:.3f}" for t in ts), f"-> best {n/min(ts)/1e6:.0f} M poses/s, median {n/sorted(ts)[len(ts)//2]/1e6:.0f}")
# raw ceiling
dc = torch.empty((n, 8), dtype=torch.float32, device="cuda"); do = torch.empty((n, 6), dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for r in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.cuda.stream(s1): dc.copy_(tc.view(n, 8), non_blocking=True)
    with torch.cuda.stream(s2): to.copy_(do, non_blocking=True)
    torch.cuda.synchronize(); t = time.perf_counter() - t0
    torch.cuda.synchronize(); t0 = time.perf_counter()
    dc.copy_(tc.view(n, 8), non_blocking=True); torch.cuda.synchronize(); ti = time.perf_counter() - t0
    t0 = time.perf_counter(); to.copy_(do, non_blocking=True); torch.cuda.synchronize(); tout = time.perf_counter() - t0
print(f"raw copies: both directions at once {t*1e3:.3f} ms; H2D alone {ti*1e3:.3f} ms ({32*n/ti/1e9:.1f} GB/s); D2H alone {tout*1e3:.3f} ms ({48*n/tout/1e9:.1f} GB/s)")
