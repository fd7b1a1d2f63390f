"""Time the UNMODIFIED reference classes (oracle/ref_import.py -> /root/reference/filters/*) next to the numpy
oracle port on the same synthetic stream, in the build container (the reference cannot travel to the GPU box).
    python tools/time_reference.py [markers] [frames]
Prints marker-updates/s of both and the port / reference factor that BASELINE.md quotes."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from aruco_slam_b200 import synth  # noqa: E402
from oracle import ekf_oracle, ref_import  # noqa: E402


def main():
    markers = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    for kind, p in (("ekf", 0.5), ("rot", 0.5)):
        s = synth.make_stream(0, markers, frames, p, kind)
        ref = ref_import.make_ekf() if kind == "ekf" else ref_import.make_rot()
        ref.observe(s.ids[0], s.poses[0])          # frame 0 adds every marker (untimed, like a warm-up)
        t0 = time.perf_counter()
        n = 0
        for t in range(1, frames):
            ref.observe(s.ids[t], s.poses[t])
            n += len(s.ids[t])
        t_ref = time.perf_counter() - t0
        o = ekf_oracle.make(kind, "as_written")
        o.observe([int(i) for i in s.ids[0]], s.poses[0])
        t0 = time.perf_counter()
        for t in range(1, frames):
            o.observe([int(i) for i in s.ids[t]], s.poses[t])
        t_port = time.perf_counter() - t0
        print(f"{kind}: {markers} markers, {frames - 1} timed frames, {n} marker-updates: reference "
              f"{n / t_ref:.1f}/s ({1e3 * t_ref / (frames - 1):.1f} ms/frame), oracle port {n / t_port:.1f}/s "
              f"({1e3 * t_port / (frames - 1):.2f} ms/frame), port/reference = {t_ref / t_port:.1f}x")


if __name__ == "__main__":
    main()
