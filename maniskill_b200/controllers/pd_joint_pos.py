"""PD joint position controller (mirror of mani_skill/agents/controllers/pd_joint_pos.py).

Parent class of the EE controllers: ``reset`` (:51-66), ``set_drive_targets`` (:68-71),
``before_simulation_step`` with interpolation (:92-98), ``get_state`` / ``set_state``
(:100-107), drive properties (:35-49).  Elementwise torch here -- the path has no IK.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence, Union

import numpy as np
import torch

from maniskill_b200 import spaces

from .base_controller import BaseController, ControllerConfig


class PDJointPosController(BaseController):
    config: "PDJointPosControllerConfig"
    _sq = None
    _tq = None
    def _stepped_by_lead(self):
        """This controller's columns were just written by the lead EE controller's launch (CombinedController fusion)."""
        self._fused_stale = True
        self._step = 0

    _fused_stale = False  # set by CombinedController when this controller's columns were written
    #                       by the fused EE launch (msik_delta_step_fused): state is gathered on demand

    def _materialise_fused(self):
        """After a fused launch the drive-target buffer holds this controller's targets and the
        qpos buffer its start positions: gather them the way set_action would have left them."""
        self._fused_stale = False
        self._sq = self.qpos
        self._tq = self.articulation.drive_targets[:, self.active_joint_indices.long()] if hasattr(
            self.articulation, "drive_targets") else self.articulation.px.cuda_articulation_target_qpos.torch()[
            self.articulation._data_index.long()][:, self.active_joint_indices.long()]

    @property
    def _start_qpos(self):
        if self._fused_stale:
            self._materialise_fused()
        return self._sq

    @_start_qpos.setter
    def _start_qpos(self, v):
        self._sq = v

    @property
    def _target_qpos(self):
        if self._fused_stale:
            self._materialise_fused()
        return self._tq

    @_target_qpos.setter
    def _target_qpos(self, v):
        self._tq = v

    def _get_joint_limits(self):
        qlimits = self.articulation.get_qlimits()[0, self.active_joint_indices.long()].cpu().numpy()
        if self.config.lower is not None:
            qlimits[:, 0] = self.config.lower
        if self.config.upper is not None:
            qlimits[:, 1] = self.config.upper
        return qlimits

    def _initialize_action_space(self):
        joint_limits = self._get_joint_limits()
        self.single_action_space = spaces.Box(joint_limits[:, 0], joint_limits[:, 1], dtype=np.float32)

    def set_drive_property(self):
        n = len(self.joints)
        stiffness = np.broadcast_to(self.config.stiffness, n)
        damping = np.broadcast_to(self.config.damping, n)
        force_limit = np.broadcast_to(self.config.force_limit, n)
        friction = np.broadcast_to(self.config.friction, n)
        for i, joint in enumerate(self.joints):
            drive_mode = self.config.drive_mode
            if not isinstance(drive_mode, str):
                drive_mode = drive_mode[i]
            joint.set_drive_properties(stiffness[i], damping[i], force_limit=force_limit[i], mode=drive_mode)
            joint.set_friction(friction[i])

    def reset(self):
        super().reset()
        self._step = 0
        mask = self.scene._reset_mask
        if self._fused_stale:
            self._materialise_fused()
        if self._start_qpos is None:
            self._start_qpos = self.qpos.clone()
        else:
            self._start_qpos[mask] = self.qpos[mask].clone()
        if self._target_qpos is None:
            self._target_qpos = self.qpos.clone()
        else:
            self._target_qpos[mask] = self.qpos[mask].clone()

    def set_drive_targets(self, targets):
        self.articulation.set_joint_drive_targets(targets, self.joints, self.active_joint_indices)

    def set_action(self, action):
        self._fused_stale = False
        action = self._preprocess_action(action)
        self._step = 0
        self._start_qpos = self.qpos
        if self.config.use_delta:
            if self.config.use_target:
                self._target_qpos = self._target_qpos + action
            else:
                self._target_qpos = self._start_qpos + action
        else:
            self._target_qpos = torch.broadcast_to(action, self._start_qpos.shape).clone()
        if self.config.interpolate:
            self._step_size = (self._target_qpos - self._start_qpos) / self._sim_steps
        else:
            self.set_drive_targets(self._target_qpos)

    def before_simulation_step(self):
        self._step += 1
        if self.config.interpolate:
            targets = self._start_qpos + self._step_size * self._step
            self.set_drive_targets(targets)

    def get_state(self) -> dict:
        if self.config.use_target:
            return {"target_qpos": self._target_qpos}
        return {}

    def set_state(self, state: dict):
        if self.config.use_target:
            self._target_qpos = state["target_qpos"]


@dataclass
class PDJointPosControllerConfig(ControllerConfig):
    lower: Union[None, float, Sequence[float]]
    upper: Union[None, float, Sequence[float]]
    stiffness: Union[float, Sequence[float]]
    damping: Union[float, Sequence[float]]
    force_limit: Union[float, Sequence[float]] = 1e10
    friction: Union[float, Sequence[float]] = 0.0
    use_delta: bool = False
    use_target: bool = False
    interpolate: bool = False
    normalize_action: bool = True
    drive_mode: Union[Sequence[str], str] = "force"
    controller_cls = PDJointPosController


class PDJointPosMimicController(PDJointPosController):
    def _get_joint_limits(self):
        joint_limits = super()._get_joint_limits()
        diff = joint_limits[0:-1] - joint_limits[1:]
        assert np.allclose(diff, 0), "Mimic joints should have the same limit"
        return joint_limits[0:1]


class PDJointPosMimicControllerConfig(PDJointPosControllerConfig):
    controller_cls = PDJointPosMimicController
