"""Trajectory text output compatible with the reference's ``outputs/trajectory_writer.py``.

One line per frame, exactly as ``TrajectoryWriter.write`` formats it (trajectory_writer.py:29-40):
``f"{timestamp / 1000:.4f} {x} {y} {z} {q0} {q1} {q2} {q3}\\n"`` where ``q0..q3 = pose[3:7]`` in STATE order
(qw first — the reference's comment says TUM ``qx qy qz qw`` but the code writes the state as it is), numbers
through ``str()`` of the numpy scalars.  Pure host code; the pose comes from ``get_poses()``."""
from __future__ import annotations

from pathlib import Path


class TrajectoryWriter:
    def __init__(self, filename: str) -> None:
        self.file = None
        self.filename = filename

    def __enter__(self):
        self.file = Path(self.filename).open("w", encoding="utf-8")
        return self

    def write(self, timestamp, pose) -> None:
        """``timestamp`` in milliseconds (cv2.CAP_PROP_POS_MSEC, main/run_slam.py:124), ``pose`` the camera pose."""
        quat = pose[3:]
        seconds = timestamp / 1000
        if self.file:
            line = f"{seconds:.4f} "
            line += f"{pose[0]} {pose[1]} {pose[2]} "
            line += f"{quat[0]} {quat[1]} {quat[2]} {quat[3]}\n"
            self.file.write(line)

    def __exit__(self, exc_type, exc_value, traceback) -> None:
        if self.file:
            self.file.close()
            self.file = None
