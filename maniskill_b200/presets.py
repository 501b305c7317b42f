"""Stock EE controller dicts of the reference's robots, built from this package's controller configs.

Each function returns the ``dict(uid -> ControllerConfig)`` that the reference agent's ``_controller_configs`` holds for
one control mode (same uids, joint lists, bounds and gains), ready for
``CombinedController(configs, articulation, control_freq, sim_freq, scene)`` -- the "class swap" integration of
INTEGRATION.md section A without touching the robot file.  bench.py, tools/ and the tests use them as the workloads.

  panda_pd_ee_delta_pose      agents/robots/panda/panda.py:103-141,192-194
  fetch_pd_ee_delta_pose      agents/robots/fetch/fetch.py:144-174,273-278
  realman_pd_ee_delta_pos     agents/robots/realman/mobile_realman.py:298-443   (the fork's robot; its one enabled EE mode)
  realman_pd_ee_delta_pose    the same dict with both arms on pd_ee_delta_pose (commented out in the fork, :443-450)
"""
from __future__ import annotations

import os
from typing import Dict

from maniskill_b200.controllers import (
    ControllerConfig,
    PDBaseForwardVelControllerConfig,
    PDBaseVelControllerConfig,
    PDEEPosControllerConfig,
    PDEEPoseControllerConfig,
    PDJointPosControllerConfig,
    PDJointPosMimicControllerConfig,
)

ASSETS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets", "robots")
PANDA_URDF = os.path.join(ASSETS, "panda_v2_kin.urdf")
FETCH_URDF = os.path.join(ASSETS, "fetch_kin.urdf")
REALMAN_URDF = os.path.join(ASSETS, "realman_dual_65B_kin.urdf")

PANDA_ARM = [f"panda_joint{i}" for i in range(1, 8)]
FETCH_ARM = ["shoulder_pan_joint", "shoulder_lift_joint", "upperarm_roll_joint", "elbow_flex_joint", "forearm_roll_joint",
             "wrist_flex_joint", "wrist_roll_joint"]
BASE_JOINTS = ["root_x_axis_joint", "root_y_axis_joint", "root_z_rotation_joint"]
REALMAN_RIGHT_ARM = [f"r_joint{i}" for i in range(1, 7)]
REALMAN_LEFT_ARM = [f"l_joint{i}" for i in range(1, 7)]
REALMAN_RIGHT_FINGERS = ["right_thumb_1_joint", "right_thumb_2_joint", "right_index_1_joint", "right_middle_1_joint",
                         "right_ring_1_joint", "right_little_1_joint"]
REALMAN_LEFT_FINGERS = [n.replace("right_", "left_") for n in REALMAN_RIGHT_FINGERS]
REALMAN_HEAD = ["head_joint1", "head_joint2"]


def _ee(cls, joints, ee_link, urdf, **kw):
    base = dict(joint_names=joints, pos_lower=-0.1, pos_upper=0.1, stiffness=1e3, damping=1e2, force_limit=100, ee_link=ee_link,
                urdf_path=urdf)
    if cls is PDEEPoseControllerConfig:
        base.update(rot_lower=-0.1, rot_upper=0.1)
    base.update(kw)
    return cls(**base)


def panda_pd_ee_delta_pose(urdf: str = PANDA_URDF) -> Dict[str, ControllerConfig]:
    return dict(
        arm=_ee(PDEEPoseControllerConfig, PANDA_ARM, "panda_hand_tcp", urdf),
        gripper=PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01, upper=0.04,
                                                stiffness=1e3, damping=1e2, force_limit=100),
    )


def fetch_pd_ee_delta_pose(urdf: str = FETCH_URDF) -> Dict[str, ControllerConfig]:
    return dict(
        arm=_ee(PDEEPoseControllerConfig, FETCH_ARM, "gripper_link", urdf),
        gripper=PDJointPosMimicControllerConfig(joint_names=["l_gripper_finger_joint", "r_gripper_finger_joint"], lower=-0.01,
                                                upper=0.05, stiffness=1e3, damping=1e2, force_limit=100),
        body=PDJointPosControllerConfig(joint_names=["head_pan_joint", "head_tilt_joint", "torso_lift_joint"], lower=-0.1, upper=0.1,
                                        stiffness=1e3, damping=1e2, force_limit=100, use_delta=True),
        base=PDBaseVelControllerConfig(joint_names=BASE_JOINTS, lower=[-1, -1, -3.14], upper=[1, 1, 3.14], damping=1000,
                                       force_limit=500),
    )


def _realman(arm_cls, urdf: str) -> Dict[str, ControllerConfig]:
    grip = lambda joints: PDJointPosControllerConfig(joint_names=joints, lower=-10.0, upper=10.0, stiffness=1e3, damping=1e2,  # noqa: E731
                                                     force_limit=100)
    return dict(
        right_arm=_ee(arm_cls, REALMAN_RIGHT_ARM, "r_link6", urdf),
        right_gripper=grip(REALMAN_RIGHT_FINGERS),
        left_arm=_ee(arm_cls, REALMAN_LEFT_ARM, "l_link6", urdf),
        left_gripper=grip(REALMAN_LEFT_FINGERS),
        # the fork passes lower = upper = None here (limits taken from the articulation); the kinematics-only fixture has
        # no head limits worth the name, so the delta range is spelled out
        body=PDJointPosControllerConfig(joint_names=REALMAN_HEAD, lower=-0.1, upper=0.1, stiffness=1e3, damping=1e2, force_limit=100,
                                        use_delta=True),
        base=PDBaseForwardVelControllerConfig(joint_names=BASE_JOINTS, lower=[-1, -3.14], upper=[1, 3.14], damping=1000,
                                              force_limit=500),
    )


def realman_pd_ee_delta_pos(urdf: str = REALMAN_URDF) -> Dict[str, ControllerConfig]:
    return _realman(PDEEPosControllerConfig, urdf)


def realman_pd_ee_delta_pose(urdf: str = REALMAN_URDF) -> Dict[str, ControllerConfig]:
    return _realman(PDEEPoseControllerConfig, urdf)
