#!/usr/bin/env python3
"""Golden vectors for the velocity controllers of SURVEY section 8 row (f)-3, produced by RUNNING THE REFERENCE'S OWN
code: ``PDJointVelController.set_action`` (pd_joint_vel.py:38-42), ``PDBaseVelController.set_action``
(pd_base_vel.py:19-33), ``PDBaseForwardVelController.set_action`` (pd_base_vel.py:50-72) with
``BaseController._preprocess_action`` / ``_clip_and_scale_action`` (base_controller.py:120-170), imported from
/root/reference with the absent third-party packages stubbed (same machinery as tools/gen_golden_controller.py).

Runs only in the build container; the output tests/golden/vel_golden.npz is committed with this script.
"""
import importlib
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import gen_golden_controller as GC  # noqa: E402

OUT = os.path.join(GC.ROOT, "tests", "golden", "vel_golden.npz")


class _VelRecorder:
    def __init__(self, qpos):
        self._qpos, self.device, self.written = qpos, qpos.device, None

    def get_qpos(self):
        return self._qpos

    def set_joint_drive_velocity_targets(self, targets, joints, joint_indices):
        self.written = (targets.clone(), joint_indices.clone())


def make(cls, config, qpos, idx, low, high, normalize=True):
    B = qpos.shape[0]
    c = object.__new__(cls)
    c.config = config
    c.scene = SimpleNamespace(num_envs=B)
    c.articulation = _VelRecorder(qpos)
    c.joints = [None] * len(idx)
    c.active_joint_indices = torch.tensor(idx)
    c._normalize_action = normalize
    c.action_space_low, c.action_space_high = torch.tensor(np.float32(low)), torch.tensor(np.float32(high))
    c.action_space = SimpleNamespace(shape=(B, len(low)))
    return c


def main() -> int:
    GC.import_reference()
    jv = importlib.import_module("mani_skill.agents.controllers.pd_joint_vel")
    bv = importlib.import_module("mani_skill.agents.controllers.pd_base_vel")
    g = torch.Generator().manual_seed(2022)
    B = 512
    out = {}
    # Fetch-like row: 15 dofs, base joints (x, y, yaw) in columns 0..2 (fetch.py:77-85,176-185)
    qpos = (torch.rand((B, 15), generator=g) * 2 - 1).float() * 3.0
    out["qpos"] = qpos.numpy()
    # pd_base_vel: lower = [-1, -1, -3.14], upper = [1, 1, 3.14] (fetch.py:176-185)
    cfg = bv.PDBaseVelControllerConfig(joint_names=None, lower=[-1, -1, -3.14], upper=[1, 1, 3.14], damping=1000, force_limit=500)
    c = make(bv.PDBaseVelController, cfg, qpos, [0, 1, 2], [-1, -1, -3.14], [1, 1, 3.14])
    a3 = (torch.rand((B, 3), generator=g) * 3 - 1.5).float()
    c.set_action(a3.clone())
    out.update(base_action=a3.numpy(), base_target_qvel=c.articulation.written[0].numpy(), base_cols=c.articulation.written[1].numpy())
    # the same without normalisation
    c = make(bv.PDBaseVelController, cfg, qpos, [0, 1, 2], [-1, -1, -3.14], [1, 1, 3.14], normalize=False)
    c.set_action(a3.clone())
    out.update(base_raw_target_qvel=c.articulation.written[0].numpy())
    # pd_base_forward_vel: 2-dim action (forward, rotation), 3 joints
    cfgf = bv.PDBaseForwardVelControllerConfig(joint_names=None, lower=[-1, -3.14], upper=[1, 3.14], damping=1000, force_limit=500)
    cf = make(bv.PDBaseForwardVelController, cfgf, qpos, [0, 1, 2], [-1, -3.14], [1, 3.14])
    a2 = (torch.rand((B, 2), generator=g) * 3 - 1.5).float()
    cf.set_action(a2.clone())
    out.update(fwd_action=a2.numpy(), fwd_target_qvel=cf.articulation.written[0].numpy())
    # pd_joint_vel on 7 arm joints (columns 4, 6, 7, 8, 9, 10, 11: arbitrary, not contiguous)
    cols = [4, 6, 7, 8, 9, 10, 11]
    cfgj = jv.PDJointVelControllerConfig(joint_names=None, lower=-1.0, upper=2.0, damping=100, force_limit=100)
    cj = make(jv.PDJointVelController, cfgj, qpos, cols, [-1.0] * 7, [2.0] * 7)
    a7 = (torch.rand((B, 7), generator=g) * 3 - 1.5).float()
    cj.set_action(a7.clone())
    out.update(joint_action=a7.numpy(), joint_target_qvel=cj.articulation.written[0].numpy(), joint_cols=cj.articulation.written[1].numpy())
    np.savez_compressed(OUT, **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", os.path.relpath(OUT), os.path.getsize(OUT), "B;", {k: np.asarray(v).shape for k, v in out.items()})
    return 0


if __name__ == "__main__":
    sys.exit(main())
