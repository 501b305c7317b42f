"""PD joint velocity controller (mirror of mani_skill/agents/controllers/pd_joint_vel.py) and the ego-centric base
controllers built on it (pd_base_vel.py) -- the velocity controllers that share a ``CombinedController`` with the EE
controller (Fetch: ``base=pd_base_vel``, agents/robots/fetch/fetch.py:176-185).

``set_action`` is ONE launch of ``msik_vel_targets`` (include/msik.h): action clip / scale, the rotation of the (x, y)
velocity by the base orientation, and the scatter into ``cuda_articulation_target_qvel`` that the reference does with
``Articulation.set_joint_drive_velocity_targets`` (utils/structs/articulation.py:743-761).  No torch fallback: a
non-CUDA articulation raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Sequence, Union

import numpy as np
import torch

from maniskill_b200 import _lib as L
from maniskill_b200 import spaces

from .base_controller import BaseController, ControllerConfig


class PDJointVelController(BaseController):
    config: "PDJointVelControllerConfig"
    _ego = L.VEL_JOINT

    def _initialize_action_space(self):
        n = len(self.joints)
        low = np.float32(np.broadcast_to(self.config.lower, n))
        high = np.float32(np.broadcast_to(self.config.upper, n))
        self.single_action_space = spaces.Box(low, high, dtype=np.float32)

    def set_drive_property(self):
        n = len(self.joints)
        damping = np.broadcast_to(self.config.damping, n)
        force_limit = np.broadcast_to(self.config.force_limit, n)
        friction = np.broadcast_to(self.config.friction, n)
        for i, joint in enumerate(self.joints):
            drive_mode = self.config.drive_mode
            if not isinstance(drive_mode, str):
                drive_mode = drive_mode[i]
            joint.set_drive_properties(0, damping[i], force_limit=force_limit[i], mode=drive_mode)
            joint.set_friction(friction[i])

    # ------------------------------------------------------------------------------------------------
    def _vel_cfg(self, action_stride: int, action_offset: int) -> "L.VelCfg":
        key = (action_stride, action_offset)
        cache = self.__dict__.setdefault("_vel_cfgs", {})
        if key in cache:
            return cache[key]
        art = self.articulation
        cols = [int(x) for x in self.active_joint_indices.tolist()]
        if len(cols) > L.MSIK_MAX_AUX:
            raise L.MsikError(f"{len(cols)} velocity-controlled joints > MSIK_MAX_AUX ({L.MSIK_MAX_AUX})")
        cfg = L.VelCfg()
        cfg.n_envs = self.scene.num_envs
        cfg.n_cols = len(cols)
        for k, c in enumerate(cols):
            cfg.col[k] = c
        cfg.ego = self._ego
        cfg.ori_col = cols[2] if self._ego != L.VEL_JOINT else 0  # pd_base_vel.py:22: "the 3rd DoF stands for orientation"
        cfg.flags = L.AUX_NORMALIZE if self._normalize_action else 0
        cfg.qpos_stride = art.px.cuda_articulation_qpos.torch().shape[1]
        cfg.qvel_stride = art.px.cuda_articulation_target_qvel.torch().shape[1]
        cfg.action_stride, cfg.action_offset = action_stride, action_offset
        if self._normalize_action:
            lo, hi = self.action_space_low, self.action_space_high
            mid = (0.5 * (hi + lo)).float().cpu().numpy()  # same fp32 roundings as _clip_and_scale_action
            half = (0.5 * (hi - lo)).float().cpu().numpy()
            for k in range(len(mid)):
                cfg.mid[k], cfg.half[k] = float(mid[k]), float(half[k])
        cache[key] = cfg
        return cfg

    def set_action_flat(self, action: torch.Tensor, start: int, end: int, aux=None):
        """``set_action(action[:, start:end])`` reading the slice in place (called by CombinedController)."""
        art = self.articulation
        if torch.device(art.device).type != "cuda" or not action.is_cuda:
            raise L.MsikError("the velocity controllers run on the GPU simulation buffers only (no CPU path)")
        B = self.scene.num_envs
        adim = self.action_space.shape[1] if B > 1 else self.action_space.shape[0]
        assert action.shape[0] == B and end - start == adim, (tuple(action.shape), start, end, adim)
        if action.dtype != torch.float32 or action.stride(-1) != 1:
            action = action.float().contiguous()
        # a (B, k) view of a wider tensor carries its own row stride and storage offset
        cfg = self._vel_cfg(action.stride(0) if B > 1 else action.shape[1], start)
        rows = None if getattr(art, "_identity_rows", False) else art._data_index
        L.check(L.load().msik_vel_targets(C.byref(cfg), L.ptr(art.px.cuda_articulation_qpos.torch()), L.ptr(rows), action.data_ptr(),
                                          L.ptr(art.px.cuda_articulation_target_qvel.torch()), torch.cuda.current_stream().cuda_stream))

    def set_action(self, action):
        action = torch.as_tensor(action, device=self.device)
        self._check_action_shape(action)
        self.set_action_flat(action, 0, action.shape[1])


@dataclass
class PDJointVelControllerConfig(ControllerConfig):
    lower: Union[float, Sequence[float]]
    upper: Union[float, Sequence[float]]
    damping: Union[float, Sequence[float]]
    force_limit: Union[float, Sequence[float]] = 1e10
    friction: Union[float, Sequence[float]] = 0.0
    normalize_action: bool = True
    drive_mode: Union[Sequence[str], str] = "force"
    controller_cls = PDJointVelController
