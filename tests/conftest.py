"""pytest configuration: the `gpu` marker and shared helpers."""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
if str(ROOT / "oracle") not in sys.path:   # tests/test_oracle_vs_reference.py imports make_golden helpers
    sys.path.append(str(ROOT / "oracle"))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def rel_err(a, b):
    """max |a-b| / max |b| (1e-9 relative in FP64)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def cov_scaled_err(a, b):
    """max |a_ij - b_ij| / sqrt(b_ii b_jj): the error of every covariance entry against ITS OWN natural
    scale (|P_ij| <= sqrt(P_ii P_jj)), so a small cross-covariance cannot hide behind max |P|."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape and a.ndim == 2, (a.shape, b.shape)
    if b.size == 0:
        return 0.0
    d = np.sqrt(np.maximum(np.diagonal(b), 1e-300))
    return float(np.max(np.abs(a - b) / np.outer(d, d)))


def cov_block_errs(a, b, cam=10):
    """relative error (max |d| / max |ref| inside the block) of the camera, cross and landmark blocks."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    out = {"cam": rel_err(a[:cam, :cam], b[:cam, :cam])}
    if b.shape[0] > cam:
        out["cross"] = rel_err(a[cam:, :cam], b[cam:, :cam])
        out["lm"] = rel_err(a[cam:, cam:], b[cam:, cam:])
    return out


PARITY_LINES = []


def assert_parity(label, state, cov, ref_state, ref_cov, tol=1e-9, floor=None):
    """The parity bar of SURVEY.md 8(c), both measures asserted: max-norm relative error of the state and of
    P, AND the per-entry scaled covariance error and the per-block (camera / cross / landmark) relative
    errors.  `floor`: the sensitivity floor of the same run (oracle with inputs x (1 + 1e-13)), reported
    beside the result.  Lines are printed in the pytest terminal summary."""
    ref_cov = np.asarray(ref_cov)
    es, ep = rel_err(state, ref_state), rel_err(cov, ref_cov)
    sc = cov_scaled_err(cov, ref_cov)
    blocks = cov_block_errs(cov, ref_cov)
    line = (f"{label}: state {es:.1e}  P max-norm {ep:.1e}  P scaled {sc:.1e}  blocks "
            + " ".join(f"{k} {v:.1e}" for k, v in blocks.items()))
    if floor is not None:
        line += f"  | sensitivity floor (inputs x (1+1e-13)): state {floor[0]:.1e} P {floor[1]:.1e}"
    PARITY_LINES.append(line)
    assert es < tol and ep < tol, line
    assert sc < tol, line
    assert all(v < tol for v in blocks.values()), line
    return es, ep, sc


def sensitivity_floor(kind, quat_mode, frames_ids, frames_poses, eps=1e-13):
    """SURVEY.md section 4 item 5 / 8(c) iv: the oracle run twice, the second time with every input scaled by
    (1 + eps); returns (state, P) max-norm relative differences = the honest floor of a free-running
    comparison."""
    from oracle import ekf_oracle
    runs = []
    for scale in (1.0, 1.0 + eps):
        o = ekf_oracle.make(kind, quat_mode)
        for ids, poses in zip(frames_ids, frames_poses):
            if len(ids):
                o.observe([int(i) for i in ids], np.asarray(poses) * scale)
        runs.append((o.state.copy(), np.asarray(o.uncertainty).copy()))
    return rel_err(runs[1][0], runs[0][0]), rel_err(runs[1][1], runs[0][1])


def pytest_terminal_summary(terminalreporter):
    if PARITY_LINES:
        terminalreporter.write_sep("-", "parity report (tolerance 1e-9 unless stated)")
        for line in PARITY_LINES:
            terminalreporter.write_line(line)


@pytest.fixture(scope="session")
def golden_steps():
    return np.load(GOLDEN / "ekf_steps.npz")


@pytest.fixture(scope="session")
def calibration():
    from aruco_slam_b200 import synth
    return synth.reference_calibration()
