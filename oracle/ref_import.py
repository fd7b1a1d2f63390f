"""Import the UNMODIFIED reference filters from /root/reference (this container only).

TEST INFRASTRUCTURE — not part of the product.  Used only by
``oracle/make_golden.py`` (to generate the committed fixtures under
``tests/golden/``) and by the optional ``tests/test_oracle_vs_reference.py``
(skipped when /root/reference is absent, i.e. on the GPU box).

Recipe: ``filters/__init__.py:8`` pulls in gtsam, and
``extended_kalman_filter.py:12`` imports a package name that does not exist in
the tree, so the modules are loaded by path under alias package names.  The
sympy lambdas (19 s / 26 s to build) are built once per process and shared by
every instance.
"""
from __future__ import annotations

import contextlib
import importlib.util
import os
import sys
import types
from pathlib import Path

REF_ROOT = Path(os.environ.get("ASB_REFERENCE_ROOT", "/root/reference"))

_cache: dict = {}


def available() -> bool:
    return (REF_ROOT / "filters" / "extended_kalman_filter.py").exists()


def _load(name: str, path: Path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@contextlib.contextmanager
def _cwd(path: Path):
    old = os.getcwd()
    os.chdir(path)
    try:
        yield
    finally:
        os.chdir(old)


def modules():
    """Return (base_filter, extended_kalman_filter, ekf_with_rotations) modules."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise RuntimeError(f"reference not present at {REF_ROOT}")
    for pkg in ("aruco_slam", "aruco_slam.filters", "filters"):
        if pkg not in sys.modules:
            m = types.ModuleType(pkg)
            m.__path__ = []
            sys.modules[pkg] = m
    base = _load("aruco_slam.filters.base_filter", REF_ROOT / "filters" / "base_filter.py")
    sys.modules["filters.base_filter"] = base
    ekf = _load("aruco_slam.filters.extended_kalman_filter",
                REF_ROOT / "filters" / "extended_kalman_filter.py")
    rot = _load("filters.ekf_with_rotations", REF_ROOT / "filters" / "ekf_with_rotations.py")
    _cache["mods"] = (base, ekf, rot)
    return _cache["mods"]


def _cached_lambdas(cls, key):
    """Build the sympy lambdas once per process, then patch initialize_h."""
    if key not in _cache:
        orig = cls.initialize_h
        probe = cls.__new__(cls)
        _cache[key] = orig(probe)
    pair = _cache[key]
    cls.initialize_h = lambda self: pair


def make_ekf(initial_pose=None, quat_mode: str = "as_written"):
    """Reference ``EKF``.  quat_mode='scalar_first' is the 2-line subclass of
    SURVEY.md §8c(i): only extended_kalman_filter.py:145-146 change."""
    import numpy as np
    _, ekf_mod, _ = modules()
    _cached_lambdas(ekf_mod.EKF, "ekf_lambdas")
    pose = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0], dtype=float) if initial_pose is None \
        else np.array(initial_pose, dtype=float)
    with _cwd(REF_ROOT):
        f = ekf_mod.EKF(pose)
    if quat_mode == "scalar_first":
        _patch_scalar_first(f, ekf_mod)
    elif quat_mode != "as_written":
        raise ValueError(quat_mode)
    return f


def _patch_scalar_first(f, ekf_mod):
    """Run the reference update with Rotation.from_quat(..., scalar_first=True)
    at the two call sites extended_kalman_filter.py:145-146 and nothing else."""
    real_rotation = ekf_mod.Rotation

    class _RotationScalarFirst:
        @staticmethod
        def from_quat(q, scalar_first=True):
            return real_rotation.from_quat(q, scalar_first=True)

    orig_update = type(f).update

    def update(self, ids, poses):
        ekf_mod.Rotation = _RotationScalarFirst
        try:
            return orig_update(self, ids, poses)
        finally:
            ekf_mod.Rotation = real_rotation

    f.update = types.MethodType(update, f)


def make_rot(initial_pose=None):
    import numpy as np
    _, _, rot_mod = modules()
    _cached_lambdas(rot_mod.EKF_Rotations, "rot_lambdas")
    pose = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0], dtype=float) if initial_pose is None \
        else np.array(initial_pose, dtype=float)
    with _cwd(REF_ROOT):
        f = rot_mod.EKF_Rotations(pose)
    return f


def base_filter_for_pnp():
    """A BaseFilter instance (for estimate_pose_of_markers only)."""
    base, _, _ = modules()
    import numpy as np
    with _cwd(REF_ROOT):
        return base.BaseFilter(np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 0], dtype=float), None)
