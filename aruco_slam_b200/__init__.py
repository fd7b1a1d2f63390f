"""aruco_slam_b200 — B200-native MEKF/EKF + IPPE hot path of supersimple33/aruco-slam.

Drop-in classes (same API as the reference's ``filters`` package):
    from aruco_slam_b200 import EKF, EKF_Rotations
backed by hand-written CUDA (sm_100a) behind the C ABI in include/aruco_slam_b200.h.
There is no CPU fallback: importing the filter classes loads libarucoslam_b200.so and
constructing a filter needs a CUDA device.
"""
from . import synth  # noqa: F401  (numpy only)

__all__ = ["EKF", "EKF_Rotations", "BaseFilter", "FilterBatch", "TrajectoryWriter", "estimate_pose_of_markers", "synth"]


def __getattr__(name):
    if name in ("EKF", "EKF_Rotations", "BaseFilter"):
        from . import filters
        return getattr(filters, name)
    if name == "FilterBatch":
        from .batch import FilterBatch
        return FilterBatch
    if name == "TrajectoryWriter":
        from .outputs import TrajectoryWriter
        return TrajectoryWriter
    if name == "estimate_pose_of_markers":
        from .ippe import estimate_pose_of_markers
        return estimate_pose_of_markers
    raise AttributeError(name)
