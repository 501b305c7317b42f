"""Shared builders for the GPU parity tests (shim articulation + controllers + oracle inputs)."""
import numpy as np
import torch

from maniskill_b200.controllers import (
    PDEEPosControllerConfig,
    PDEEPoseControllerConfig,
)
from maniskill_b200.shim import FakeArticulation
from maniskill_b200.structs.pose import Pose
from oracle import oracle as O


def sample_q(frames, B, seed=0, shrink=0.1, rest=None, noise=0.02):
    rng = np.random.default_rng(seed)
    if rest is not None:
        return np.asarray(rest)[None] + rng.normal(0, noise, (B, len(rest)))
    lim = np.clip(O.joint_limits(frames), -3.0, 3.0)
    lo = lim[:, 0] + shrink * (lim[:, 1] - lim[:, 0])
    hi = lim[:, 1] - shrink * (lim[:, 1] - lim[:, 0])
    return rng.uniform(lo, hi, (B, len(lim)))


PANDA_REST = [0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4]  # agents/robots/panda/panda.py:35-48


def make_art(robot, B, **kw):
    urdf, ee, _ = robot
    return FakeArticulation(urdf, B, ee_link_names=[ee], **kw)


def fill_qpos(art, frames, anc_idx, q_anc, seed=0):
    """Write chain joint values into the PhysX qpos buffer (other dofs random), return fp32-rounded q_anc."""
    B = art.num_envs
    rng = np.random.default_rng(seed + 77)
    rows = rng.uniform(0.0, 0.04, (B, art.max_dof)).astype(np.float32)
    rows[:, anc_idx] = q_anc.astype(np.float32)
    buf = art.px.cuda_articulation_qpos.torch()
    buf.normal_()  # rows of other articulations: garbage that must not be touched / read
    buf[art._data_index.long()] = torch.from_numpy(rows).cuda()
    return rows[:, anc_idx].astype(np.float64), rows


def pose_cfg(robot, **kw):
    urdf, ee, joints = robot
    base = dict(joint_names=joints, pos_lower=-0.1, pos_upper=0.1, rot_lower=-0.1, rot_upper=0.1, stiffness=1e3, damping=1e2,
                force_limit=100, ee_link=ee, urdf_path=urdf)
    base.update(kw)
    return PDEEPoseControllerConfig(**base)


def pos_cfg(robot, **kw):
    urdf, ee, joints = robot
    base = dict(joint_names=joints, pos_lower=-0.1, pos_upper=0.1, stiffness=1e3, damping=1e2, force_limit=100, ee_link=ee,
                urdf_path=urdf)
    base.update(kw)
    return PDEEPosControllerConfig(**base)


def set_link_poses_from_fk(art, ctrl, root_pose_np=None):
    """Stand-in for PhysX: root pose given (default identity), EE world pose = root * FK(qpos)."""
    B = art.num_envs
    if root_pose_np is None:
        root_pose_np = np.tile([0, 0, 0, 1.0, 0, 0, 0], (B, 1))
    root = Pose.create(torch.from_numpy(root_pose_np.astype(np.float32)).cuda())
    art.set_root_pose(root)
    fk = ctrl.kinematics.forward_kinematics(art.get_qpos())
    art.set_link_pose(ctrl.ee_link.name, root * fk)


def cond_strata(err, cond, edges=((0, 50), (50, 200), (200, 1e3), (1e3, 1e12))):
    out = {}
    for lo, hi in edges:
        m = (cond >= lo) & (cond < hi)
        if m.any():
            out[(lo, hi)] = (int(m.sum()), float(err[m].max()), float(np.median(err[m])))
    return out
