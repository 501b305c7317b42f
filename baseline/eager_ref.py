"""Eager-PyTorch restatement of the reference's GPU controller path (TIMED BASELINE).

The reference runs this path through ``pytorch_kinematics==0.7.5`` (setup.py:46), which
is absent here and not installable (no network), so the stock code cannot be executed.
This file restates, op for op in eager torch, what the reference executes per
``set_action`` on GPU sim:

  PDEEPoseController._clip_and_scale_action      pd_ee_pose.py:197-210
  BaseController.qpos / Articulation.get_qpos    base_controller.py:101-103, articulation.py:613-622
  Kinematics.compute_ik (use_delta_ik_solver)    kinematics.py:142-143,156-172
      pk_chain.jacobian(q0): python loop over the chain frames building 4x4 transforms,
      Jacobian in the EE frame, then a second FK pass and a 6x6 block rotation to the
      base frame (pytorch_kinematics jacobian.calc_jacobian, recalled structure)
      torch.linalg.pinv(J) @ action
  Articulation.set_joint_drive_targets           articulation.py:731-738

Used only by ``bench.py`` (reference-GPU timing, labelled "restated") and tests (as a
second opinion next to the fp64 oracle).  Never imported by maniskill_b200/.
"""
from __future__ import annotations

import math
from typing import List, Sequence

import torch

from maniskill_b200.urdf import UrdfJoint, parse_urdf, rpy_matrix, serial_chain


class EagerChain:
    """Per-frame constants as torch tensors (what pk keeps per Frame)."""

    def __init__(self, urdf_path: str, end_link: str, device, dtype=torch.float32):
        _, joints = parse_urdf(urdf_path)
        self.frames: List[UrdfJoint] = serial_chain(joints, end_link)
        self.device, self.dtype = device, dtype
        self.offsets = []
        self.axes = []
        for f in self.frames:
            T = torch.eye(4, dtype=torch.float64)
            T[:3, :3] = torch.from_numpy(rpy_matrix(f.rpy))
            T[:3, 3] = torch.from_numpy(f.xyz)
            self.offsets.append(T.to(device=device, dtype=dtype))
            self.axes.append(torch.tensor(f.axis / max(1e-30, float((f.axis**2).sum()) ** 0.5), dtype=dtype, device=device))
        self.n = sum(1 for f in self.frames if f.movable)

    # axis-angle -> 4x4, batched (pk Frame.get_transform)
    def _joint_tf(self, i: int, th: torch.Tensor) -> torch.Tensor:
        f = self.frames[i]
        N = th.shape[0]
        T = torch.eye(4, dtype=self.dtype, device=self.device).repeat(N, 1, 1)
        a = self.axes[i]
        if f.type in ("revolute", "continuous"):
            c, s = torch.cos(th), torch.sin(th)
            C = 1 - c
            x, y, z = a[0], a[1], a[2]
            R = torch.stack(
                [
                    c + x * x * C, x * y * C - z * s, x * z * C + y * s,
                    y * x * C + z * s, c + y * y * C, y * z * C - x * s,
                    z * x * C - y * s, z * y * C + x * s, c + z * z * C,
                ],
                -1,
            ).reshape(N, 3, 3)
            T[:, :3, :3] = R
        elif f.type == "prismatic":
            T[:, :3, 3] = a[None, :] * th[:, None]
        return self.offsets[i][None] @ T

    def forward_kinematics(self, th: torch.Tensor) -> torch.Tensor:
        N = th.shape[0]
        T = torch.eye(4, dtype=self.dtype, device=self.device).repeat(N, 1, 1)
        k = 0
        for i, f in enumerate(self.frames):
            if f.movable:
                T = T @ self._joint_tf(i, th[:, k])
                k += 1
            else:
                T = T @ self.offsets[i][None]
        return T

    def jacobian(self, th: torch.Tensor, ret_eef_pose: bool = False):
        """(N, 6, n) base-frame Jacobian, rows [linear; angular]."""
        N, ndof = th.shape
        j_eef = torch.zeros((N, 6, ndof), dtype=self.dtype, device=self.device)
        cur = torch.eye(4, dtype=self.dtype, device=self.device).repeat(N, 1, 1)
        cnt = 0
        for i in reversed(range(len(self.frames))):
            f = self.frames[i]
            if f.type in ("revolute", "continuous"):
                cnt += 1
                Rt = cur[:, :3, :3].transpose(1, 2)
                axis_in_eef = Rt @ self.axes[i]
                pos = (Rt @ cur[:, :3, 3].unsqueeze(2)).squeeze(2)
                j_eef[:, :, -cnt] = torch.cat((torch.cross(axis_in_eef, pos, dim=1), axis_in_eef), dim=-1)
            elif f.type == "prismatic":
                cnt += 1
                j_eef[:, :3, -cnt] = (self.axes[i].repeat(N, 1, 1) @ cur[:, :3, :3])[:, 0, :]
            if f.movable:
                cur = self._joint_tf(i, th[:, -cnt]) @ cur
            else:
                cur = self.offsets[i][None] @ cur
        pose = self.forward_kinematics(th)
        rot = pose[:, :3, :3]
        j_tr = torch.zeros((N, 6, 6), dtype=self.dtype, device=self.device)
        j_tr[:, :3, :3] = rot
        j_tr[:, 3:, 3:] = rot
        j_w = j_tr @ j_eef
        if ret_eef_pose:
            return j_w, pose
        return j_w


def clip_and_scale_pose_action(action, low, high, rot_lower):
    """pd_ee_pose.py:197-210 + gym_utils.py:102-105, eager."""
    pos = torch.clip(action[:, :3], -1, 1)
    pos = 0.5 * (high[:3] + low[:3]) + 0.5 * (high[:3] - low[:3]) * pos
    rot = action[:, 3:].clone()
    nrm = torch.linalg.norm(rot, axis=1)
    rot[nrm > 1] = torch.mul(rot, 1 / nrm[:, None])[nrm > 1]
    rot = rot * rot_lower
    return torch.hstack([pos, rot])


class EagerDeltaController:
    """pd_ee_delta_pose set_action on GPU sim, eager torch, PhysX-layout buffers."""

    def __init__(self, urdf_path, end_link, qpos_buf, target_buf, data_index, anc_idx: Sequence[int],
                 ctrl_idx: Sequence[int], pos_bound=0.1, rot_lower=-0.1):
        dev = qpos_buf.device
        self.chain = EagerChain(urdf_path, end_link, dev)
        self.qpos_buf, self.target_buf, self.data_index = qpos_buf, target_buf, data_index.long()
        self.anc = list(anc_idx)
        self.ctrl = torch.tensor(list(ctrl_idx), device=dev, dtype=torch.long)
        self.ctrl_in_anc = [self.anc.index(int(c)) for c in ctrl_idx]
        self.low = torch.tensor([-pos_bound] * 3 + [rot_lower] * 3, device=dev)
        self.high = torch.tensor([pos_bound] * 3 + [-rot_lower] * 3, device=dev)
        self.rot_lower = rot_lower
        self.max_dof = qpos_buf.shape[1]
        self.gx, self.gy = torch.meshgrid(self.data_index, self.ctrl, indexing="ij")

    def set_action(self, action: torch.Tensor):
        a = clip_and_scale_pose_action(action, self.low, self.high, self.rot_lower)
        qpos = self.qpos_buf[self.data_index, : self.max_dof]  # articulation.get_qpos()
        _start = qpos[..., self.ctrl]  # controller.qpos (_start_qpos)
        q0 = qpos[:, self.anc]
        J = self.chain.jacobian(q0)
        if len(self.ctrl_in_anc) != len(self.anc):
            J = J[:, :, self.ctrl_in_anc]  # intended masking
            q0 = q0[:, self.ctrl_in_anc]
        dq = torch.linalg.pinv(J) @ a.unsqueeze(-1)
        tq = q0 + dq.squeeze(-1)
        self.target_buf[self.gx, self.gy] = tq
        return tq
