#!/usr/bin/env python
"""bench.py — MEKF marker-updates/s on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm (oracle port)

Headline workload (config.workload "C2"): BASELINE.json configs[1] — 4096 independent
synthetic MEKF sequences x 32 markers x 1000 frames PER GPU (weak scaling: sequences
shard across ranks with no data-path collective).  One "step" = one pass of the hot path
(EKF.observe for every frame of every sequence) over that batch from freshly constructed
filters.  `value` = marker-updates/s with the observations already resident in HBM;
`e2e` = the same through the public Python API with host buffers (H2D of the step's
observations from pinned memory + D2H of the resulting camera poses inside the timed
region).  Extra workloads measured in the same run and reported under "workloads":
C1 (one sequence, per-frame observe() latency, both classes), C2 with EKF_Rotations (N = 330, wide engine),
C3 (one EKF, 2000-marker map, N = 6010, the covariance sweep), C4 (1M-corner batched IPPE) and C5
(32768 sequences x 64 markers x 1000 frames, p = 0.25, both classes, total over the ranks).
See DESIGN.md §Measurement for the byte/flop accounting.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

# FP64 flops of k_ippe_square per corner set: (2 x 536.18 M dfma + 499.39 M dmul + 333.70 M dadd) / 1 M sets, ncu
# smsp__sass_thread_inst_executed_op_d{fma,mul,add}_pred_on.sum (profiles/r02_k_ippe.txt)
IPPE_FLOPS_PER_SET = 1905.45
METRIC = "mekf_marker_updates_per_sec"
UNIT = "marker-updates/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--seqs", type=int, default=4096, help="sequences per GPU (C2)")
    ap.add_argument("--markers", type=int, default=32)
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--p-visible", type=float, default=0.5)
    ap.add_argument("--unique", type=int, default=64,
                    help="distinct synthetic sequences generated per rank and tiled to --seqs")
    ap.add_argument("--kind", default="ekf", choices=["ekf", "rot"])
    ap.add_argument("--quat-mode", default="as_written", choices=["as_written", "scalar_first"])
    ap.add_argument("--no-extras", action="store_true", help="skip the C3 / C4 legs")
    ap.add_argument("--c5-seqs", type=int, default=32768, help="C5 sequences in total (sharded over the ranks)")
    ap.add_argument("--leg-budget-s", type=float, default=170.0,
                    help="wall-clock budget of one extra batched leg; a leg that would exceed it stops at a "
                         "whole number of 100-frame slices and says so (frames_run)")
    ap.add_argument("--only", default=None, choices=["c1", "c2rot", "c3", "c4", "c3s", "c5ekf", "c5rot"],
                    help="profiling aid: run just that leg and print its JSON")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU baseline budget")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = max(mx, float(r[2]))
                for name, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def flops_batch_form(n_state: int, k: int, rows: int, width: int) -> float:
    """Algorithmic FP64 flops of one frame: P update, gain solve,
    Cholesky, PH^T, S, dx for m = rows*k measurement rows."""
    m = rows * k
    n = n_state
    return 2.0 * n * n * m + 2.0 * n * m * m + m ** 3 / 3.0 + 2.0 * width * n * m \
        + 2.0 * width * m * m + 2.0 * n * m


def flops_executed(n_state: int, k: int, rows: int, width: int) -> float:
    """FP64 flops the kernels actually execute for one frame: the covariance update on the TRIANGLE (N^2 m instead
    of the contract's dense 2 N^2 m) and the triangular gain solve (N m^2 instead of 2 N m^2); the rest as in
    flops_batch_form.  Reported beside the contract figure so that `frac` cannot be read as pipe utilisation."""
    m = rows * k
    n = n_state
    return float(n) * n * m + float(n) * m * m + m ** 3 / 3.0 + 2.0 * width * n * m + 2.0 * width * m * m + 2.0 * n * m


def pinned_like(a: np.ndarray):
    import torch
    t = torch.empty(a.shape, dtype=torch.from_numpy(np.empty(0, a.dtype)).dtype, pin_memory=True)
    out = t.numpy()
    out[...] = a
    return t, out


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's algorithm on the host cores
# ------------------------------------------------------------------------------------------
def _cpu_worker(job):
    seq, markers, frames, p_vis, kind, quat_mode = job
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from aruco_slam_b200 import synth
    from oracle import ekf_oracle
    s = synth.make_stream(seq, markers, frames, p_vis, kind)
    o = ekf_oracle.make(kind, quat_mode)
    t0 = time.perf_counter()
    n = 0
    for t in range(frames):
        o.observe([int(i) for i in s.ids[t]], s.poses[t])
        n += len(s.ids[t])
    return n, time.perf_counter() - t0


def cpu_port_throughput(args, workers: int, frames: int, reps: int = 1):
    """marker-updates/s of the numpy oracle: `workers` processes, one sequence each."""
    import multiprocessing as mp
    # one BLAS thread per worker process: the children read this when they import numpy
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = "1"
    jobs = [(s, args.markers, frames, args.p_visible, args.kind, args.quat_mode)
            for s in range(workers * reps)]
    t0 = time.perf_counter()
    if workers == 1:
        res = [_cpu_worker(j) for j in jobs]
    else:
        with mp.get_context("spawn").Pool(workers) as pool:
            pool.map(_cpu_worker, jobs[:workers])  # warm: imports, page-in
            t0 = time.perf_counter()
            res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    total = sum(r[0] for r in res)
    return total / wall, total, wall


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 64))
    frames = 500
    # one step = one 500-frame sequence per worker (~0.5 s at ~1 ms / frame / sequence)
    vals = []
    for _ in range(args.warmup):
        pass  # the pool's own warm pass (imports, page-in) is inside cpu_port_throughput
    t_all = time.perf_counter()
    for _ in range(max(1, args.steps)):
        v, total, wall = cpu_port_throughput(args, workers, frames, reps=1)
        vals.append(v)
        if time.perf_counter() - t_all > 150:
            break
    value = float(np.mean(vals))
    sample = (f"{workers} sequences x {args.markers} markers x {frames} frames per step, one "
              f"process per sequence, numpy oracle port of the reference algorithm "
              f"(dense (I-KH)P as written)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(vals), "warmup": args.warmup, "ms_per_step": 1e3 * wall, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(json.dumps(line))


def workload_config(args):
    return {"workload": f"C2: {args.seqs} independent MEKF sequences x {args.markers} markers x "
                        f"{args.frames} frames per GPU (BASELINE.json configs[1])",
            "filter": "EKF (filters/extended_kalman_filter.py)" if args.kind == "ekf"
            else "EKF_Rotations (filters/ekf_with_rotations.py)",
            "quat_mode": args.quat_mode, "sequences_per_gpu": args.seqs, "markers": args.markers,
            "frames": args.frames, "p_visible": args.p_visible,
            "distinct_streams_per_gpu": min(args.unique, args.seqs),
            "streams_note": f"{min(args.unique, args.seqs)} distinct synthetic sequences per rank tiled to "
                            f"{args.seqs} (every sequence runs its own filter and the whole 3.4 GB stream is "
                            f"read); value_unique = one step with {args.seqs} distinct sequences",
            "parallelism": "sequence shards, no data-path collective, final camera-pose gather",
            "cache": "inputs larger than L2 (3.4 GB of observations per step)"}


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def timed(torch, dist, fn, steps):
    """EXACTLY `steps` calls bracketed by barrier + synchronize; CUDA events on the current
    stream; max over ranks."""
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def leg_c3(torch, stream_ptr, hbm_peak, fp64_peak):
    """C3: one EKF with a 2000-marker map (N = 6010, P = 289 MB), 20 markers per frame."""
    import ctypes as C
    from aruco_slam_b200 import EKF, _lib, synth
    km, dist = synth.reference_calibration()
    timed_frames, warm = 200, 20
    stream, n_build = synth.make_map_build_stream(2000, 20, warm + timed_frames)
    f = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dist,
            quat_mode="as_written", max_landmarks=2000)
    f.set_stream(stream_ptr)
    ids, poses, off = stream.packed()
    f.stream_upload(ids, poses, off)
    f.stream_run(0, n_build + warm)  # untimed map build + warm-up frames
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    f.stream_run(n_build + warm, n_build + warm + timed_frames)
    e1.record()
    torch.cuda.synchronize()
    f.synchronize()
    ms = e0.elapsed_time(e1)
    n = 6010
    sec = ms * 1e-3
    # Bytes the sweep has to move as built: it keeps the LOWER TRIANGLE only (one read + one write = 8 N^2);
    # the dense-equivalent contract figure of SURVEY.md 8(d) (16 N^2: what a full-matrix update would move) is
    # reported beside it under its own name, not as the achieved bandwidth.
    ach = 8.0 * n * n * timed_frames / sec / 1e9
    dense_eq = 16.0 * n * n * timed_frames / sec / 1e9
    tf = float(n) * n * 60 * timed_frames / sec / 1e12   # N^2 m flops per frame on the triangle (m = 60)
    # e2e: per-frame observe() with host buffers + camera pose read-back every frame
    t_frames = 50
    extra, _ = synth.make_map_build_stream(2000, 20, t_frames, seed=77)
    nb = 2000 // 20
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(nb, nb + t_frames):
        f.observe(extra.ids[t], extra.poses[t])
        cam, _ = f.get_poses()
    wall = time.perf_counter() - t0
    del f
    return {
        "workload": "C3: single EKF, 2000-marker map (N=6010, FP64 P 289 MB), 20 markers/frame",
        "value": 20.0 * timed_frames / (ms * 1e-3), "unit": UNIT, "us_per_frame": 1e3 * ms / timed_frames,
        "frames": timed_frames, "gpu_launches_per_frame": 3,
        "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                     "frac": ach / hbm_peak, "traffic": 256.3e6,
                     "dense_equivalent": {"achieved": dense_eq, "frac": dense_eq / hbm_peak,
                                          "note": "16 N^2 B per frame (SURVEY.md 8d contract figure: one read + one "
                                                  "write of a DENSE P) / whole-frame time"},
                     "note": "achieved = 8 N^2 B per frame (lower triangle read + write, what k_sweep moves) / "
                             "whole-frame time (k_f2 + k_sweep on the critical path, k_f1 of the next frame "
                             "overlapped); traffic = dram read + write bytes of one k_sweep launch, ncu --set full "
                             "(profiles/r02_k_sweep.txt: 154.35 MB read + 101.95 MB written).  The kernel is bound by the FP64 tensor pipe, see "
                             "roofline_fp64"},
        "roofline_fp64": {"bound": "tensor", "achieved": tf, "peak": fp64_peak, "unit": "TFLOP/s",
                          "frac": tf / fp64_peak, "peak_source": "asb_fp64_peak(which=1): mma.sync m8n8k4 f64 (DMMA) "
                          "chains timed live with CUDA events; MEASURED_PEAKS.json has no FP64 entry",
                          "note": "N^2 m flops per frame (triangle) / whole-frame time"},
        "e2e": {"value": 20.0 * t_frames / wall, "unit": UNIT, "h2d_bytes_per_step": 20 * 52,
                "d2h_bytes_per_step": 8 * 6010},
    }


def leg_c3_sharded(torch, dist, rank, world, local, stream_ptr, hbm_peak):
    """C3 with the covariance rows sharded over `world` GPUs (filter_shard.cu): every rank feeds the
    same stream; two NCCL collectives per frame inside the library; max over ranks."""
    from aruco_slam_b200 import synth
    from aruco_slam_b200.sharded import ShardedFilter
    timed_frames, warm = 200, 20
    stream, n_build = synth.make_map_build_stream(2000, 20, warm + timed_frames)
    f = ShardedFilter("ekf", np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), 2000, dist=dist, rank=rank, world=world,
                      device=local, quat_mode="scalar_first")
    f.set_stream(stream_ptr)
    ids, poses, off = stream.packed()
    f.stream_upload(ids, poses, off)
    f.stream_run(0, n_build + warm)
    ms = timed(torch, dist, lambda: f.stream_run(n_build + warm, n_build + warm + timed_frames), 1)
    f.synchronize()
    cam, _ = f.get_poses()
    n = 6010
    ach = 8.0 * n * n * timed_frames / (ms * 1e-3) / 1e9
    # driver-visible multi-GPU correctness: the same frames on the single-GPU engine (rank 0), state and this
    # rank's covariance rows compared with the sharded result (max-norm relative error)
    parity = {}
    try:
        rows = f.local_rows()
        if rank == 0:
            from aruco_slam_b200 import EKF
            km, dc = synth.reference_calibration()
            g1 = EKF(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]), camera_matrix=km, dist_coeffs=dc,
                     quat_mode="scalar_first", max_landmarks=2000)
            g1.stream_upload(ids, poses, off)
            g1.stream_run(0, n_build + warm + timed_frames)
            st1, p1 = g1.state, g1.uncertainty
            del g1
            parity = {"parity_rel_state": float(np.max(np.abs(f.state - st1)) / np.max(np.abs(st1))),
                      "parity_rel_cov_rows": float(np.max(np.abs(rows - p1[f.row0:f.row0 + rows.shape[0]]))
                                                   / np.max(np.abs(p1))),
                      "parity_note": "row-sharded engine vs the single-GPU engine on the same 320 frames (rank 0's "
                                     "rows), free-running, scalar-first quaternion mode (the as-written reset is "
                                     "chaotic, SURVEY.md F3); tolerance 1e-9"}
    except Exception as exc:
        parity = {"parity_error": repr(exc)}
    out = {"workload": f"C3 row-sharded: single EKF, 2000-marker map (N=6010), 20 markers/frame, covariance rows "
                       f"over {world} GPU(s), NCCL all-reduce of P[T,T] + all-gather of the V panel per frame",
           "value": 20.0 * timed_frames / (ms * 1e-3), "unit": UNIT, "us_per_frame": 1e3 * ms / timed_frames,
           "n_gpus": world, "frames": timed_frames, "rows_per_gpu": f.row1 - f.row0, "quat_mode": "scalar_first",
           "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak * world, "unit": "GB/s",
                        "frac": ach / (hbm_peak * world), "traffic": None,
                        "note": "8 N^2 B per frame (the bytes of P every rank set together has to move once) / frame "
                                "time against world x the measured HBM peak; the "
                                "rectangular row sweep does 2x the FP64 work of the triangle sweep and each "
                                "frame pays two NCCL collectives"},
           "camera_finite": bool(np.isfinite(cam).all())}
    out.update(parity)
    f.close()
    return out


def leg_c4(torch, stream_ptr, hbm_peak, fp64_peak):
    """C4: batched IPPE over 1M corner sets."""
    from aruco_slam_b200 import estimate_pose_of_markers, synth
    from aruco_slam_b200.ippe import estimate_pose_device
    km, dist = synth.reference_calibration()
    n = 1_000_000
    base = synth.make_corner_sets(125_000, km, dist, seed=11)
    corners = np.ascontiguousarray(np.tile(base, (8, 1, 1)))
    d_c = torch.from_numpy(corners).cuda()
    d_o = torch.empty((n, 6), dtype=torch.float64, device="cuda")
    for _ in range(3):
        estimate_pose_device(d_c.data_ptr(), n, km, dist, 0.16, d_o.data_ptr(), stream=stream_ptr)
    torch.cuda.synchronize()
    reps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        estimate_pose_device(d_c.data_ptr(), n, km, dist, 0.16, d_o.data_ptr(), stream=stream_ptr)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # e2e through the public host API with page-locked host buffers (corners in, poses out)
    keep_c, h_corners = pinned_like(corners)
    keep_o, h_out = pinned_like(np.zeros((n, 6)))
    estimate_pose_of_markers(h_corners, km, dist, 0.16, out=h_out)   # warm (buffers of the call grow once)
    walls = []
    for _ in range(5):   # median of five calls (each: 32 MB in, kernels, 48 MB out, synchronous return)
        t0 = time.perf_counter()
        out = estimate_pose_of_markers(h_corners, km, dist, 0.16, out=h_out)
        walls.append(time.perf_counter() - t0)
    wall = sorted(walls)[2]
    ach = 80.0 * n / (ms * 1e-3) / 1e9
    return {"workload": "C4: batched IPPE square-marker pose, 1M corner sets", "value": n / (ms * 1e-3),
            "unit": "poses/s", "ms_per_batch": ms,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                         "frac": ach / hbm_peak, "traffic": None,
                         "note": "80 B per set (32 in + 48 out); the kernel is FP64-pipe bound, see roofline_fp64"},
            "roofline_fp64": {"bound": "fp64", "achieved": IPPE_FLOPS_PER_SET * n / (ms * 1e-3) / 1e12, "peak": fp64_peak,
                              "unit": "TFLOP/s", "frac": IPPE_FLOPS_PER_SET * n / (ms * 1e-3) / 1e12 / fp64_peak,
                              "flops_per_set": IPPE_FLOPS_PER_SET, "fp64_pipe_active_pct_ncu": 63.7,
                              "note": "flops per corner set counted by ncu on this kernel (2 x dfma + dmul + dadd "
                                      "thread instructions, profiles/r02_k_ippe.txt); peak = DFMA peak measured live "
                                      "(asb_fp64_peak); a third of the issued FP64 instructions are DMUL / DADD, "
                                      "which occupy the pipe for one flop"},
            "e2e": {"value": n / wall, "unit": "poses/s", "h2d_bytes_per_step": 32 * n,
                    "d2h_bytes_per_step": 48 * n}, "finite": bool(np.isfinite(out).all())}


def leg_c1(torch, stream_ptr):
    """C1 (BASELINE.json configs[0]): ONE sequence, 10 static markers, 1000 frames, p = 0.8 — the latency
    case: per-frame observe() through the drop-in classes with host buffers and the camera pose read back
    every frame (what main/run_slam.py does), both classes, next to the numpy oracle port on one host core."""
    from aruco_slam_b200 import EKF, EKF_Rotations, synth
    from oracle import ekf_oracle
    km, dist = synth.reference_calibration()
    pose0 = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
    out = {"workload": "C1: 1 sequence x 10 markers x 1000 frames, p = 0.8, per-frame observe() + get_poses() "
                       "with host buffers (BASELINE.json configs[0])"}
    for kind, cls in (("ekf", EKF), ("rot", EKF_Rotations)):
        s = synth.make_stream(0, 10, 1000, 0.8, kind)
        kw = dict(camera_matrix=km, dist_coeffs=dist, max_landmarks=16)
        if kind == "ekf":
            kw["quat_mode"] = "as_written"
        for rep in range(2):          # first pass warms the kernels / staging buffers
            f = cls(pose0, **kw)
            f.set_stream(stream_ptr)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for t in range(s.n_frames):
                f.observe(s.ids[t], s.poses[t])
                cam, _ = f.get_poses()
            wall = time.perf_counter() - t0
            del f
        o = ekf_oracle.make(kind, "as_written")
        t0 = time.perf_counter()
        for t in range(200):
            o.observe([int(i) for i in s.ids[t]], s.poses[t])
        cpu = sum(len(i) for i in s.ids[:200]) / (time.perf_counter() - t0)
        out[kind] = {"value": s.n_updates / wall, "unit": UNIT, "us_per_frame": 1e6 * wall / s.n_frames,
                     "frames": s.n_frames, "marker_updates": s.n_updates, "camera_finite": bool(np.isfinite(cam).all()),
                     "cpu_port_value": cpu, "cpu_port_sample": "first 200 frames, oracle/ekf_oracle.py, 1 core",
                     "e2e": {"value": s.n_updates / wall, "unit": UNIT,
                             "h2d_bytes_per_step": 52 * s.n_updates // s.n_frames, "d2h_bytes_per_step": 8 * 10 * 11}}
    return out


def leg_batch(torch, dist, world, rank, local, stream_ptr, args, name, kind, seqs_total, markers, frames, p_vis,
              fp64_peak, fp64_src, unique=64, full_upload=False):
    """A batched workload next to the headline: `seqs_total` sequences IN TOTAL, sharded over the ranks
    (no data-path collective), one step = every frame of every sequence from freshly constructed filters.
    The host holds `unique` distinct streams (sequence s reads stream s % unique: 32768 x 64 x 1000 distinct
    observations would be 25 GB); every sequence runs its own filter.  Time: CUDA events, max over ranks."""
    from aruco_slam_b200 import FilterBatch, synth
    n_seq = max(1, seqs_total // world)
    u = min(unique, n_seq)
    while n_seq % u:
        u -= 1
    ids, poses, off = synth.make_batch(u, markers, frames, p_vis, kind, seq_base=5000 + rank * u, unique=u)
    rows, width, d = (3, 13, 3) if kind == "ekf" else (7, 20, 10)
    n_state = 10 + d * markers
    counts = np.diff(off).reshape(u, frames)
    quat_mode = args.quat_mode if kind == "ekf" else "scalar_first"
    b = FilterBatch(kind, n_seq, markers, markers, quat_mode=quat_mode, device=local)
    b.set_stream(stream_ptr)
    if full_upload:
        ids_f, poses_f, off_f = synth.make_batch(n_seq, markers, frames, p_vis, kind, seq_base=5000 + rank * u, unique=u)
        keep = [pinned_like(a) for a in (ids_f, poses_f, off_f)]
        up = lambda: b.upload(keep[0][1], keep[1][1], keep[2][1])            # noqa: E731
        h2d = ids_f.nbytes + poses_f.nbytes + off_f.nbytes
    else:
        keep = [pinned_like(a) for a in (ids, poses, off)]
        up = lambda: b.upload(keep[0][1], keep[1][1], keep[2][1], n_streams=u)   # noqa: E731
        h2d = ids.nbytes + poses.nbytes + off.nbytes
    up()
    b.run(0, min(10, frames))      # warm-up: every kernel of the path, same shapes
    b.synchronize()
    b.reset()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall = time.perf_counter()
    e0.record()
    done = 0
    while done < frames:
        nxt = min(frames, done + 100)
        b.run(done, nxt)
        done = nxt
        if done < frames:
            b.synchronize()
            # stop at a whole slice if the rest would not fit the leg's budget (reported as frames_run)
            if (time.perf_counter() - t_wall) * (1.0 + 100.0 / done) > args.leg_budget_s:
                break
    e1.record()
    torch.cuda.synchronize()
    ms = float(e0.elapsed_time(e1))
    cams, status = b.cameras()
    rep = n_seq // u
    upd = float(counts[:, :done].sum()) * rep
    kc_hist = list(zip(*np.unique(counts[:, :done], return_counts=True)))
    flops = float(sum(flops_batch_form(n_state, int(k), rows, width) * c for k, c in kc_hist)) * rep
    flops_x = float(sum(flops_executed(n_state, int(k), rows, width) * c for k, c in kc_hist)) * rep
    tot = torch.tensor([upd, flops, flops_x], dtype=torch.float64, device="cuda")
    mx = torch.tensor([ms, -float(done)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tot)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    upd_all, flops_all, ms_max, frames_min = float(tot[0]), float(tot[1]), float(mx[0]), int(-float(mx[1]))
    tf = flops_all / (ms_max * 1e-3) / 1e12
    tf_x = float(tot[2]) / (ms_max * 1e-3) / 1e12
    leg = {"workload": name, "value": upd_all / (ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
           "sequences_total": n_seq * world, "sequences_per_gpu": n_seq, "markers": markers, "frames": frames,
           "frames_run": frames_min, "p_visible": p_vis, "n_state": n_state, "engine": b.engine,
           "filter": "EKF" if kind == "ekf" else "EKF_Rotations", "quat_mode": quat_mode, "ms_per_step": ms_max,
           "distinct_streams_per_gpu": u, "scaling": "strong (total sequences fixed, sharded over the ranks)",
           "roofline": {"bound": "fp64", "achieved": tf, "peak": fp64_peak * world, "unit": "TFLOP/s",
                        "frac": tf / (fp64_peak * world), "traffic": None, "peak_source": fp64_src,
                        "executed": {"achieved": tf_x, "frac": tf_x / (fp64_peak * world),
                                     "note": "flops the kernels execute (triangle update N^2 m, triangular solve "
                                             "N m^2): the FP64-pipe utilisation to compare with ncu"},
                        "note": "algorithmic batch-form flops F(N,m,k) of SURVEY.md 8d / time against world x the "
                                "FP64 peak measured live; the covariance update counts 2 N^2 m although the "
                                "kernels compute its triangle (N^2 m)"},
           "valid": bool(np.isfinite(cams).all() and not status.any())}
    if done == frames and ms < 20e3:
        # e2e through the public API: host buffers in, camera poses out, every step
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        up()
        b.reset()
        b.run(0, frames)
        cams, status = b.cameras()
        wall = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(wall, op=dist.ReduceOp.MAX)
        leg["e2e"] = {"value": upd_all / float(wall.item()), "unit": UNIT, "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": cams.nbytes + status.nbytes,
                      "note": "full-size host stream" if full_upload else
                              f"host stream = the {u} distinct sequences (asb_batch_upload_tiled)"}
    else:
        leg["e2e"] = None
    del b
    return leg



def run_b200(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    import ctypes as C
    from aruco_slam_b200 import FilterBatch, _lib, synth

    lib = _lib.lib()
    hbm_peak, peak_src = measured_peaks()
    fp64 = C.c_double()
    _lib.check(lib.asb_fp64_peak(local, 0, C.byref(fp64)))
    fp64_peak = fp64.value
    _lib.check(lib.asb_fp64_peak(local, 1, C.byref(fp64)))
    dmma_peak = fp64.value
    fp64_src = ("asb_fp64_peak: dependent-free DFMA chains (FMA pipe) / mma.sync m8n8k4 f64 chains (DMMA) timed "
                "live on this GPU with CUDA events — one shared FP64 pipe; MEASURED_PEAKS.json has no FP64 entry")
    # a non-default stream: the library launches on it and torch.cuda.Event times it
    side = torch.cuda.Stream()
    torch.cuda.set_stream(side)
    stream_ptr = torch.cuda.current_stream().cuda_stream
    assert stream_ptr != 0

    def extra_leg(which):
        if which == "c1":
            return leg_c1(torch, stream_ptr)
        if which == "c2rot":
            return leg_batch(torch, dist, world, rank, local, stream_ptr, args,
                             f"C2-ROT: {args.seqs} EKF_Rotations sequences x 32 markers x 1000 frames per GPU "
                             f"(N = 330, BASELINE.json configs[1] secondary)", "rot", args.seqs * world, 32, 1000,
                             0.5, max(fp64_peak, dmma_peak), fp64_src, full_upload=True)
        if which in ("c5ekf", "c5rot"):
            kind = "ekf" if which == "c5ekf" else "rot"
            return leg_batch(torch, dist, world, rank, local, stream_ptr, args,
                             f"C5-{kind.upper()}: {args.c5_seqs} sequences x 64 markers x 1000 frames in total, "
                             f"p = 0.25 (BASELINE.json configs[4])", kind, args.c5_seqs, 64, 1000, 0.25,
                             max(fp64_peak, dmma_peak), fp64_src)
        if which == "c3s":
            return leg_c3_sharded(torch, dist, rank, world, local, stream_ptr, hbm_peak)
        if which == "c3":
            return leg_c3(torch, stream_ptr, hbm_peak, dmma_peak)
        return leg_c4(torch, stream_ptr, hbm_peak, fp64_peak)

    if args.only:
        leg = extra_leg(args.only)
        if rank == 0:
            emit(json.dumps(leg))
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- synthetic observations of this rank's shard --------------------------------------
    t_gen = time.perf_counter()
    ids, poses, off = synth.make_batch(args.seqs, args.markers, args.frames, args.p_visible, args.kind,
                                       seq_base=rank * args.unique, unique=args.unique)
    n_updates = int(off[-1])
    rows, width, d = (3, 13, 3) if args.kind == "ekf" else (7, 20, 10)
    n_state = 10 + d * args.markers
    counts = np.diff(off)
    flops = float(sum(flops_batch_form(n_state, int(k), rows, width) * c
                      for k, c in zip(*np.unique(counts, return_counts=True))))
    flops_x = float(sum(flops_executed(n_state, int(k), rows, width) * c
                        for k, c in zip(*np.unique(counts, return_counts=True))))
    gen_s = time.perf_counter() - t_gen
    keep_i, h_ids = pinned_like(ids)
    keep_p, h_poses = pinned_like(poses)
    keep_o, h_off = pinned_like(off)

    batch = FilterBatch(args.kind, args.seqs, args.markers, args.markers, quat_mode=args.quat_mode,
                        device=local)
    batch.set_stream(stream_ptr)
    batch.upload(h_ids, h_poses, h_off)
    torch.cuda.synchronize()

    def step_resident():
        batch.reset()
        batch.run(0, args.frames)

    from aruco_slam_b200 import dist as asb_dist

    def step_e2e():
        batch.upload(h_ids, h_poses, h_off)
        batch.reset()
        batch.run(0, args.frames)
        cams_, status_ = batch.cameras()
        # the only inter-GPU step of the sharded path: final gather of the camera poses to rank 0
        asb_dist.gather_results(cams_, world * args.seqs, dist, device=torch.device("cuda", local))
        return cams_, status_

    for _ in range(args.warmup):
        step_resident()
    with ClockSampler(local) as clk:
        ms = timed(torch, dist, step_resident, args.steps)
    cams, status = batch.cameras()
    ok = bool(np.isfinite(cams).all() and not status.any())
    ms_step = ms / args.steps
    value = world * n_updates / (ms_step * 1e-3)

    # e2e through the public API with host buffers
    step_e2e()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    wall = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(wall, op=dist.ReduceOp.MAX)
    e2e_value = world * n_updates * args.steps / float(wall.item())
    h2d = ids.nbytes + poses.nbytes + off.nbytes
    d2h = cams.nbytes + status.nbytes

    ach_tf = flops / (ms_step * 1e-3) / 1e12
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64",
        "data": f"synthetic (SURVEY.md 8d generator; {args.unique} distinct sequences per rank tiled to "
                f"{args.seqs})",
        "config": workload_config(args),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": 2 * args.steps,
        "clocks": clk.summary(),
        "roofline": {"bound": "fp64", "achieved": ach_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": ach_tf / fp64_peak, "traffic": None, "peak_source": fp64_src,
                     "executed": {"achieved": flops_x / (ms_step * 1e-3) / 1e12,
                                  "frac": flops_x / (ms_step * 1e-3) / 1e12 / fp64_peak,
                                  "note": "flops k_batch executes (triangle update, triangular solve): compare with "
                                          "ncu sm__pipe_fp64_cycles_active (profiles/r02_k_batch.txt: 25.9 %)"},
                     "note": "k_batch keeps each filter in shared memory; HBM traffic is the observation "
                             "stream only, so the FP64 FMA pipe bounds it. achieved = algorithmic "
                             "batch-form flops F(N,m,k) of SURVEY.md 8d / time; peak = DFMA peak measured "
                             "live by asb_fp64_peak",
                     "hbm_gbs": (h2d + 2 * args.seqs * 8 * (n_state + n_state * (n_state + 1) // 2))
                     / (ms_step * 1e-3) / 1e9, "hbm_peak": hbm_peak, "hbm_peak_source": peak_src},
        "valid": ok, "marker_updates_per_step_per_gpu": n_updates, "gen_seconds": gen_s,
    }
    del batch
    del keep_i, keep_p, keep_o, h_ids, h_poses, h_off
    if not args.no_extras:
        line["workloads"] = {}
        # batched legs that shard over the ranks (collective-free, every rank takes part)
        names = ["c5ekf", "c5rot"] + (["c3s"] if world > 1 else ["c2rot"])
        for which in names:
            try:
                leg = extra_leg(which)
            except Exception as exc:  # keep the headline line even if an extra leg fails
                leg = {"error": repr(exc)}
            line["workloads"][{"c5ekf": "c5_ekf", "c5rot": "c5_rot", "c3s": "c3_sharded", "c2rot": "c2_rot"}[which]] = leg
    if rank == 0 and world == 1:
        if not args.no_extras:
            for name in ("c1", "c3", "c4"):
                try:
                    line["workloads"][name] = extra_leg(name)
                except Exception as exc:
                    line["workloads"][name] = {"error": repr(exc)}
            # the HBM-side kernel of the path (north_star's 70 % target) next to the headline roofline
            if "roofline" in line["workloads"].get("c3", {}):
                line["roofline_hbm_kernel"] = dict(line["workloads"]["c3"]["roofline"], kernel="k_sweep (C3 leg)")
            # the same step on args.seqs DISTINCT sequences: the number must not move
            try:
                cores = max(1, min(os.cpu_count() or 1, 32))
                ids_u, poses_u, off_u = synth.make_batch(args.seqs, args.markers, args.frames, args.p_visible, args.kind,
                                                         seq_base=20000, unique=args.seqs, workers=cores)
                bu = FilterBatch(args.kind, args.seqs, args.markers, args.markers, quat_mode=args.quat_mode, device=local)
                bu.set_stream(stream_ptr)
                bu.upload(ids_u, poses_u, off_u)
                bu.run(0, args.frames)
                bu.synchronize()

                def step_u():
                    bu.reset()
                    bu.run(0, args.frames)
                ms_u = timed(torch, None, step_u, 2) / 2
                cu, su = bu.cameras()
                line["value_unique"] = {"value": float(off_u[-1]) / (ms_u * 1e-3), "unit": UNIT, "ms_per_step": ms_u,
                                        "distinct_streams": args.seqs, "steps": 2,
                                        "valid": bool(np.isfinite(cu).all() and not su.any())}
                del bu, ids_u, poses_u, off_u
            except Exception as exc:
                line["value_unique"] = {"error": repr(exc)}
        # bounded sample: whole sequences of the same workload on ONE core until the budget is spent
        frames = min(args.frames, 1000)
        total, wall_cpu, nseq_cpu = 0, 0.0, 0
        try:
            from threadpoolctl import threadpool_limits
            one_thread = threadpool_limits(limits=1)   # "cores": 1 means one BLAS thread as well
        except ImportError:
            one_thread = None
        while wall_cpu < args.cpu_seconds and nseq_cpu < 64:
            n_upd, dt = _cpu_worker((nseq_cpu, args.markers, frames, args.p_visible, args.kind, args.quat_mode))
            total += n_upd
            wall_cpu += dt
            nseq_cpu += 1
        if one_thread is not None:
            one_thread.restore_original_limits()
        line["cpu_baseline"] = {
            "value": total / wall_cpu, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{nseq_cpu} sequences x {args.markers} markers x {frames} frames of the same workload, "
                      f"oracle/ekf_oracle.py (numpy restatement of the reference algorithm), "
                      f"{wall_cpu:.1f} s on {os.cpu_count()} host cores (1 used)"}
    if rank == 0:
        emit(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


_RESULT_FD = None


def emit(text):
    """The result line goes to the process's ORIGINAL stdout; everything else printed while the run is going on
    (NCCL's version banner, library chatter) lands on stderr, so stdout carries the one JSON line only."""
    sys.stderr.flush()
    if _RESULT_FD is None:
        print(text, flush=True)
    else:
        os.write(_RESULT_FD, (text + "\n").encode())


def main():
    global _RESULT_FD
    args = parse_args()
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
