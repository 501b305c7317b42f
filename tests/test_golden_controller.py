"""The oracle against golden vectors produced by RUNNING THE REFERENCE'S OWN controller code
(tools/gen_golden_controller.py: pd_ee_pose.py / pd_joint_pos.py / kinematics.py GPU delta branch /
pose.py / gym_utils.py / rotation_conversions.py imported from the reference, fp32, with
pytorch_kinematics' Jacobian -- the only absent piece on this path -- served by the oracle).
The reference computes in fp32, the oracle in fp64: tolerances are fp32 rounding, not slack."""
import os

import numpy as np
import pytest

from conftest import FETCH, PANDA
from oracle import oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "controller_golden.npz"))
F64 = lambda k: G[k].astype(np.float64)  # noqa: E731


@pytest.fixture(scope="module")
def panda():
    return O.load_chain(PANDA[0], PANDA[1])


def test_ee_pose_at_base_matches_reference_pose_algebra():
    eb = O.ee_pose_at_base(F64("panda_root"), F64("panda_ee_world"))  # Pose.inv / Pose.__mul__, pose.py:187-220
    assert np.abs(eb - G["panda_ee_pose_at_base"]).max() < 2e-6


def test_set_action_pose_matches_reference(panda):
    ref = O.controller_step(panda, G["panda_qpos"], G["panda_action"], anc_idx=list(range(7)), ctrl_in_anc=list(range(7)),
                            ee_pose_base=F64("panda_ee_pose_at_base"))
    assert np.abs(ref["action"] - G["panda_pose_scaled_action"]).max() < 1e-7  # _clip_and_scale_action incl. the sign flip
    assert np.abs(ref["target_pose"] - G["panda_pose_target_pose"]).max() < 5e-7  # compute_target_pose
    # q0 + torch.linalg.pinv(J) @ action as executed by the reference (fp32)
    assert G["panda_pose_cols"].tolist() == list(range(7))
    J, _, _ = O.jacobian(panda, F64("panda_qpos")[:, :7])
    assert np.linalg.cond(J).max() < 20
    assert np.abs(ref["target_qpos"] - G["panda_pose_target_qpos"]).max() < 3e-6


def test_set_action_pos_only_matches_reference(panda):
    ref = O.controller_step(panda, G["panda_qpos"], G["panda_action"][:, :3], anc_idx=list(range(7)), ctrl_in_anc=list(range(7)),
                            pos_only=True, ee_pose_base=F64("panda_ee_pose_at_base"))
    assert np.abs(ref["target_pose"] - G["panda_pos_target_pose"]).max() < 5e-7
    assert np.abs(ref["target_qpos"] - G["panda_pos_target_qpos"]).max() < 3e-6


@pytest.mark.parametrize("frame", ["root_translation:root_aligned_body_rotation", "root_translation:body_aligned_body_rotation",
                                   "body_translation:root_aligned_body_rotation", "body_translation:body_aligned_body_rotation"])
def test_compute_target_pose_frames(frame):
    t = O.compute_target_pose(F64("prev_pose"), F64("panda_pose_scaled_action"), True, frame)
    assert np.abs(t - G["target_pose_" + frame.replace(":", "__")]).max() < 5e-7


def test_compute_target_pose_absolute_and_pos_controllers():
    t = O.compute_target_pose(None, F64("abs_action"), False, "root_translation:root_aligned_body_rotation")
    assert np.abs(t - G["target_pose_absolute"]).max() < 5e-7
    for f in ("root_translation", "body_translation"):
        t = O.compute_target_pose(F64("prev_pose"), F64("panda_pose_scaled_action")[:, :3], True, f, pos_only=True)
        assert np.abs(t - G["target_pos_" + f]).max() < 2e-7


def test_gripper_mimic_controller_is_bit_exact():
    a = np.clip(G["gripper_action"], -1, 1).astype(np.float32)
    lo, hi = np.float32(-0.01), np.float32(0.04)
    v = np.float32(0.5) * (hi + lo) + np.float32(0.5) * (hi - lo) * a  # gym_utils.py:102-105 in fp32
    assert np.array_equal(np.broadcast_to(v, (a.shape[0], 2)), G["gripper_target_qpos"])


def test_fetch_masked_delta_matches_reference_expression():
    fr = O.load_chain(FETCH[0], FETCH[1])
    mask = np.zeros(11, bool)
    mask[4:] = True
    ref = O.ik_delta_step(fr, F64("fetch_q"), F64("fetch_action"), col_mask=mask)
    J, _, _ = O.jacobian(fr, F64("fetch_q"))
    cond = np.linalg.cond(J[:, :, mask])
    d = np.abs(ref - G["fetch_target_qpos"]).max(1)
    assert (cond < 50).sum() > 300
    assert d[cond < 50].max() < 2e-5  # fp32 pinv of the reference itself: error ~ cond * eps * |dq|
