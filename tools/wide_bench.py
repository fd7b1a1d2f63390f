"""Timing of the batched engines on a B200: marker-updates/s for a (kind, n_seq, n_markers, n_frames, p) shape.
usage: python tools/wide_bench.py rot 4096 32 100 0.5 [engine] [unique]"""
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from aruco_slam_b200 import FilterBatch, synth  # noqa: E402

kind, n_seq, n_markers, n_frames, p = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
engine = sys.argv[6] if len(sys.argv) > 6 else "auto"
unique = int(sys.argv[7]) if len(sys.argv) > 7 else 64
ids, poses, off = synth.make_batch(n_seq, n_markers, n_frames, p, kind, unique=min(unique, n_seq))
b = FilterBatch(kind, n_seq, n_markers, n_markers, quat_mode="scalar_first", engine=engine)
b.upload(ids, poses, off)
b.synchronize()
best = 1e9
for rep in range(3):
    b.reset()
    b.synchronize()
    t0 = time.perf_counter()
    b.run()
    b.synchronize()
    best = min(best, time.perf_counter() - t0)
cams, st = b.cameras()
print(f"{kind} engine={b.engine} n_seq={n_seq} markers={n_markers} frames={n_frames} p={p}: "
      f"{best * 1e3:.1f} ms, {len(ids) / best / 1e6:.2f} M marker-updates/s, "
      f"{best / n_frames * 1e6:.1f} us/frame, finite={np.isfinite(cams).all()} status_nonzero={int((st != 0).sum())}")
