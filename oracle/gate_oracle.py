"""CPU restatement of ``BaseFilter.correct_poses`` (/root/reference/filters/base_filter.py:274-325), the
orientation-ambiguity gate.  TEST INFRASTRUCTURE (never imported by the product).  Literal: scipy
``Rotation`` calls in the reference's order; ``QUAT_THRESHOLD = 50`` (base_filter.py:32).  parity unpinned by
the reference itself: the function is never called there and has no test; pinned here by
tests/test_oracle_vs_reference.py against the reference's own method on a live instance."""
from __future__ import annotations

import numpy as np
from scipy.spatial.transform import Rotation

QUAT_THRESHOLD = 50
CAM_DIMS, LM_DIMS = 10, 10


def angle_deg(state, index, pose):
    rot_mc = Rotation.from_quat(state[3:7], scalar_first=True)                       # BASE:293-294
    lm_i = LM_DIMS * index + CAM_DIMS                                               # BASE:301
    state_rot_ml = Rotation.from_quat(state[lm_i + 3:lm_i + 7], scalar_first=True)  # BASE:304-305
    rot_cl = Rotation.from_euler("xyz", pose[3:])                                   # BASE:308
    r_diff = (rot_mc * rot_cl) * state_rot_ml.inv()                                 # BASE:311-314
    angle = np.degrees(r_diff.magnitude())                                          # BASE:316
    return min(angle, 360 - angle)                                                  # BASE:317


def correct_poses(state, landmarks, ids, poses, alt_poses=None, threshold=QUAT_THRESHOLD):
    """(corrected_ids, corrected_poses, choice) — choice[o] 0 dropped, 1 first solution, 2 alternative."""
    out_ids, out_poses, choice = [], [], []
    for o, (idx, pose) in enumerate(zip(ids, poses)):
        if int(idx) not in landmarks:                                               # BASE:297-298
            choice.append(0)
            continue
        index = landmarks[int(idx)]
        if angle_deg(state, index, pose) < threshold:                               # BASE:319-321
            out_ids.append(int(idx)); out_poses.append(pose); choice.append(1)
        elif alt_poses is not None and angle_deg(state, index, alt_poses[o]) < threshold:
            out_ids.append(int(idx)); out_poses.append(alt_poses[o]); choice.append(2)
        else:
            choice.append(0)
    return out_ids, out_poses, choice
