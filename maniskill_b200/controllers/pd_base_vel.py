"""Ego-centric base velocity controllers (mirror of mani_skill/agents/controllers/pd_base_vel.py): the (x, y) velocity
of the action is rotated by the base orientation (qpos of the controller's 3rd joint) inside ``msik_vel_targets``."""
from __future__ import annotations

import numpy as np

from maniskill_b200 import _lib as L
from maniskill_b200 import spaces

from .pd_joint_vel import PDJointVelController, PDJointVelControllerConfig


class PDBaseVelController(PDJointVelController):
    """pd_base_vel.py:10-33."""

    _ego = L.VEL_BASE

    def _initialize_action_space(self):
        # At least support xy-plane translation and z-axis rotation
        assert len(self.joints) >= 3, len(self.joints)
        super()._initialize_action_space()


class PDBaseVelControllerConfig(PDJointVelControllerConfig):
    controller_cls = PDBaseVelController


class PDBaseForwardVelController(PDJointVelController):
    """pd_base_vel.py:40-72: 2-dim action (forward velocity, rotation velocity) for 3 joints."""

    _ego = L.VEL_FORWARD

    def _initialize_action_space(self):
        assert len(self.joints) >= 3, len(self.joints)
        low = np.float32(np.broadcast_to(self.config.lower, 2))
        high = np.float32(np.broadcast_to(self.config.upper, 2))
        self.single_action_space = spaces.Box(low, high, dtype=np.float32)


class PDBaseForwardVelControllerConfig(PDJointVelControllerConfig):
    controller_cls = PDBaseForwardVelController
