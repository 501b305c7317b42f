"""Batched arithmetic of the reference's trajectory action conversion
(mani_skill/trajectory/utils/actions/conversion.py:52-81,100-183): turning pd_joint_pos
demonstrations into pd_ee_delta_pose / pd_ee_delta_pos actions.  The reference does this one
step and one env at a time with Pinocchio FK + numpy / sapien.Pose ("TODO: parallelize / use GPU",
conversion.py:68); here the FK is ``msik_fk`` and the pose -> action mapping ``msik_pose_to_delta_action``,
both over all envs at once.  Stepping the two simulators (conversion.py:130,206) stays with the caller.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from maniskill_b200 import _lib as L
from maniskill_b200.kinematics import _stream
from maniskill_b200.structs.pose import Pose


def target_ee_pose_from_qpos(arm_controller, full_qpos: torch.Tensor, arm_target_qpos: Optional[torch.Tensor] = None) -> Pose:
    """conversion.py:136-143: world pose of the EE link for ``full_qpos`` with the arm joints replaced by the
    original controller's ``_target_qpos`` (root pose * FK; ``pin_model.compute_forward_kinematics`` there)."""
    q = full_qpos.clone()
    if arm_target_qpos is not None:
        q[:, arm_controller.active_joint_indices.long()] = arm_target_qpos
    return arm_controller.kinematics.forward_kinematics(q, root_pose=arm_controller.articulation.pose)


def delta_pose_to_pd_ee_delta(arm_controller, target_ee_pose: Pose, ee_pose: Optional[Pose] = None,
                              pos_only: Optional[bool] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """conversion.py:60-81 + 152-183 for every env: normalised arm action that moves ``ee_pose`` (default: the
    controller's current EE link pose) to ``target_ee_pose``, and the mask of envs whose action was NOT clipped
    (the reference retries the others, conversion.py:168-178).  Both poses in the world frame, as in the reference."""
    conf = arm_controller.config
    assert conf.use_delta and conf.normalize_action, "conversion targets normalised delta controllers (conversion.py:70-71)"
    if pos_only is None:
        pos_only = conf.frame == "root_translation"
    cur = Pose.create(arm_controller.ee_link.pose if ee_pose is None else ee_pose, device=arm_controller.device)
    tgt = Pose.create(target_ee_pose, device=arm_controller.device)
    t, c = tgt.raw_pose.contiguous().float(), cur.raw_pose.contiguous().float()
    B = t.shape[0]
    assert c.shape == (B, 7), (c.shape, t.shape)
    low, high = arm_controller.action_space_low.float().cpu(), arm_controller.action_space_high.float().cpu()
    nd = 3 if pos_only else 6
    out = torch.empty((B, nd), dtype=torch.float32, device=t.device)
    exact = torch.empty((B,), dtype=torch.uint8, device=t.device)
    lo3 = (C.c_float * 3)(*[float(x) for x in low[:3]])
    hi3 = (C.c_float * 3)(*[float(x) for x in high[:3]])
    rl, rh = (float(low[3]), float(high[3])) if low.numel() >= 6 else (-1.0, 1.0)
    L.check(L.load().msik_pose_to_delta_action(lo3, hi3, rl, rh, 1 if pos_only else 0, t.data_ptr(), c.data_ptr(), B,
                                               out.data_ptr(), exact.data_ptr(), _stream()))
    return out, exact.bool()
