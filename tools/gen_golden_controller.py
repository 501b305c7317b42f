#!/usr/bin/env python3
"""Golden vectors produced by RUNNING THE REFERENCE'S OWN controller code.

Imports ``mani_skill/agents/controllers/pd_ee_pose.py`` (and with it ``base_controller.py``,
``pd_joint_pos.py``, ``utils/kinematics.py``, ``utils/structs/pose.py``, ``utils/gym_utils.py``,
``utils/geometry/rotation_conversions.py``) straight from /root/reference and executes

  * ``PDEEPoseController.set_action`` / ``PDEEPosController.set_action``  (pd_ee_pose.py:91-115)
    end to end: ``_preprocess_action`` -> ``_clip_and_scale_action`` -> ``ee_pose_at_base`` ->
    ``compute_target_pose`` -> ``Kinematics.compute_ik`` (GPU delta branch, kinematics.py:142-172,
    including the reference's ``torch.linalg.pinv``) -> ``set_drive_targets``;
  * ``PDJointPosMimicController.set_action`` (pd_joint_pos.py:73-90) for the gripper columns;
  * ``compute_target_pose`` for every frame / use_delta combination.

What cannot be imported is stubbed, and ONLY that: the absent third-party packages (sapien,
gymnasium, pytorch_kinematics, ...) become inert stub modules, and the one object the reference
obtains from pytorch_kinematics on this path -- ``pk_chain.jacobian(q)`` -- is served by the fp64
oracle (oracle/oracle.py::jacobian, itself checked against finite differences), cast to fp32 like the
reference's tensors.  Everything else that runs is reference code, in fp32 as in GPU sim.

Runs only in the build container (/root/reference is absent on the GPU box); the output
tests/golden/controller_golden.npz is committed with this script.
"""
import importlib
import importlib.abc
import importlib.machinery
import logging
import os
import sys
import types
from pathlib import Path
from types import SimpleNamespace
from unittest.mock import MagicMock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..")
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden", "controller_golden.npz")
ABSENT = {"sapien", "gymnasium", "pytorch_kinematics", "fast_kinematics", "transforms3d", "trimesh", "dacite", "h5py", "tyro",
          "lxml", "cv2", "PIL", "tqdm", "IPython", "imageio", "matplotlib"}


class _Stub(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name != "__path__":
            raise AttributeError(name)
        m = MagicMock(name=f"{self.__name__}.{name}")
        setattr(self, name, m)
        return m


class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path, target=None):
        if fullname.split(".")[0] in ABSENT:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)

    def create_module(self, spec):
        m = _Stub(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


def import_reference():
    """The reference's controller modules, without executing the heavy package __init__ files
    (mani_skill/__init__ imports every env, which needs SAPIEN)."""
    sys.meta_path.insert(0, _Finder())
    sys.path.insert(0, REF)
    for name in ("mani_skill", "mani_skill.agents", "mani_skill.envs"):
        m = types.ModuleType(name)
        m.__path__ = [os.path.join(REF, *name.split("."))]
        sys.modules[name] = m
    ms = sys.modules["mani_skill"]
    ms.logger = logging.getLogger("mani_skill")
    ms.PACKAGE_ASSET_DIR = Path(REF) / "mani_skill" / "assets"
    ms.ASSET_DIR = Path("/tmp")
    ms.format_path = lambda p: p
    import sapien  # the stub

    sapien.Pose = type("Pose", (), {})  # pose.py does isinstance(x, sapien.Pose)
    ee = importlib.import_module("mani_skill.agents.controllers.pd_ee_pose")
    jp = importlib.import_module("mani_skill.agents.controllers.pd_joint_pos")
    kin = importlib.import_module("mani_skill.agents.controllers.utils.kinematics")
    pose = importlib.import_module("mani_skill.utils.structs.pose")
    return ee, jp, kin, pose


class _Recorder:
    """articulation stand-in: hands out qpos / root pose, records set_joint_drive_targets."""

    def __init__(self, qpos, root_pose):
        self._qpos, self.pose, self.device = qpos, root_pose, qpos.device
        self.written = None

    def get_qpos(self):
        return self._qpos

    def set_joint_drive_targets(self, targets, joints, joint_indices):
        self.written = (targets.clone(), joint_indices.clone())


def make_ee_controller(ee, kinmod, posemod, cls, config, frames, qpos, root_pose, ee_pose, anc_idx, ctrl_idx, normalize):
    """A reference controller object with its collaborators faked and its own methods intact."""
    from oracle import oracle as O

    B = qpos.shape[0]
    c = object.__new__(cls)
    c.config = config
    c.scene = SimpleNamespace(num_envs=B, _reset_mask=torch.ones(B, dtype=torch.bool))
    c.articulation = _Recorder(qpos, root_pose)
    c.joints = [None] * len(ctrl_idx)
    c.active_joint_indices = torch.tensor(ctrl_idx, dtype=torch.int32).long()
    c.ee_link = SimpleNamespace(pose=ee_pose)
    c._sim_steps = 5
    c._target_pose = None
    c._normalize_action = normalize
    nd = 3 if cls is ee.PDEEPosController else 6
    low = np.float32([config.pos_lower] * 3 + ([config.rot_lower] * 3 if nd == 6 else []))
    high = np.float32([config.pos_upper] * 3 + ([config.rot_upper] * 3 if nd == 6 else []))
    c.action_space_low, c.action_space_high = torch.from_numpy(low), torch.from_numpy(high)
    c.action_space = SimpleNamespace(shape=(B, nd))
    k = object.__new__(kinmod.Kinematics)  # the reference's Kinematics, GPU branch
    k.use_gpu_ik = True
    k.device = qpos.device
    k.active_ancestor_joint_idxs = list(anc_idx)

    def jacobian(q):  # the ONE thing pytorch_kinematics would have supplied
        J, _, _ = O.jacobian(frames, q.double().numpy())
        return torch.from_numpy(J).to(q.dtype)

    k.pk_chain = SimpleNamespace(jacobian=jacobian)
    c.kinematics = k
    return c


def main() -> int:
    ee, jp, kinmod, posemod = import_reference()
    from oracle import oracle as O

    Pose = posemod.Pose
    assets = os.path.join(ROOT, "maniskill_b200", "assets", "robots")
    out = {}
    g = torch.Generator().manual_seed(2022)
    B = 512
    # ---------------- Panda, pd_ee_delta_pose / pd_ee_delta_pos through set_action -----------------
    frames = O.load_chain(os.path.join(assets, "panda_v2_kin.urdf"), "panda_hand_tcp")
    rest = torch.tensor([0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4, 0.04, 0.04])
    qpos = (rest[None] + 0.02 * torch.randn((B, 9), generator=g)).float()
    qpos[B // 2:, :7] += 0.5 * (torch.rand((B - B // 2, 7), generator=g) - 0.5)  # wider spread for half of them
    _, R, p = O.jacobian(frames, qpos[:, :7].double().numpy())
    fk = torch.from_numpy(np.concatenate([p, O.matrix_to_quaternion(R)], 1)).float()
    rq = torch.randn((B, 4), generator=g)
    rq = rq / rq.norm(dim=1, keepdim=True)
    root = Pose.create_from_pq(torch.randn((B, 3), generator=g).float(), rq.float())
    eew = root * Pose.create(fk)  # world pose of the EE link as PhysX would report it
    action = (torch.rand((B, 6), generator=g) * 3 - 1.5).float()  # exercises clip and the unit-norm rule
    cfg = ee.PDEEPoseControllerConfig(joint_names=None, pos_lower=-0.1, pos_upper=0.1, rot_lower=-0.1, rot_upper=0.1,
                                      stiffness=1e3, damping=1e2, force_limit=100, ee_link="panda_hand_tcp", urdf_path="")
    c = make_ee_controller(ee, kinmod, posemod, ee.PDEEPoseController, cfg, frames, qpos, root, eew, range(7), range(7), True)
    c.set_action(action.clone())
    tq, cols = c.articulation.written
    out.update(panda_qpos=qpos.numpy(), panda_root=root.raw_pose.numpy(), panda_ee_world=eew.raw_pose.numpy(),
               panda_action=action.numpy(), panda_pose_target_qpos=tq.numpy(), panda_pose_cols=cols.numpy(),
               panda_pose_target_pose=c._target_pose.raw_pose.numpy(),
               panda_pose_scaled_action=c._clip_and_scale_action(action.clone()).numpy(),
               panda_ee_pose_at_base=c.ee_pose_at_base.raw_pose.numpy())
    cfg3 = ee.PDEEPosControllerConfig(joint_names=None, pos_lower=-0.1, pos_upper=0.1, stiffness=1e3, damping=1e2, force_limit=100,
                                      ee_link="panda_hand_tcp", urdf_path="")
    c3 = make_ee_controller(ee, kinmod, posemod, ee.PDEEPosController, cfg3, frames, qpos, root, eew, range(7), range(7), True)
    c3.set_action(action[:, :3].clone())
    out.update(panda_pos_target_qpos=c3.articulation.written[0].numpy(), panda_pos_target_pose=c3._target_pose.raw_pose.numpy())
    # ---------------- compute_target_pose for every frame / mode -----------------------------------
    prev = Pose.create(fk)
    a_scaled = c._clip_and_scale_action(action.clone())
    for frame in ("root_translation:root_aligned_body_rotation", "root_translation:body_aligned_body_rotation",
                  "body_translation:root_aligned_body_rotation", "body_translation:body_aligned_body_rotation"):
        cf = ee.PDEEPoseControllerConfig(joint_names=None, pos_lower=-0.1, pos_upper=0.1, rot_lower=-0.1, rot_upper=0.1,
                                         stiffness=1e3, damping=1e2, frame=frame)
        cc = object.__new__(ee.PDEEPoseController)
        cc.config = cf
        out["target_pose_" + frame.replace(":", "__")] = cc.compute_target_pose(prev, a_scaled).raw_pose.numpy()
    cf = ee.PDEEPoseControllerConfig(joint_names=None, pos_lower=-0.1, pos_upper=0.1, rot_lower=-0.1, rot_upper=0.1, stiffness=1e3,
                                     damping=1e2, use_delta=False)
    cc = object.__new__(ee.PDEEPoseController)
    cc.config = cf
    abs_action = torch.cat([fk[:, :3], (torch.rand((B, 3), generator=g) * 2 - 1) * 3.0], 1).float()
    out.update(abs_action=abs_action.numpy(), target_pose_absolute=cc.compute_target_pose(prev, abs_action).raw_pose.numpy())
    for frame in ("root_translation", "body_translation"):
        cf = ee.PDEEPosControllerConfig(joint_names=None, pos_lower=-0.1, pos_upper=0.1, stiffness=1e3, damping=1e2, frame=frame)
        cc = object.__new__(ee.PDEEPosController)
        cc.config = cf
        out["target_pos_" + frame] = cc.compute_target_pose(prev, a_scaled[:, :3]).raw_pose.numpy()
    out["prev_pose"] = fk.numpy()
    # ---------------- gripper: PDJointPosMimicController.set_action --------------------------------
    gcfg = jp.PDJointPosMimicControllerConfig(joint_names=None, lower=-0.01, upper=0.04, stiffness=1e3, damping=1e2, force_limit=100)
    gc = object.__new__(jp.PDJointPosMimicController)
    gc.config = gcfg
    gc.scene = SimpleNamespace(num_envs=B)
    gc.articulation = _Recorder(qpos, root)
    gc.joints = [None, None]
    gc.active_joint_indices = torch.tensor([7, 8])
    gc._normalize_action = True
    gc.action_space_low, gc.action_space_high = torch.tensor([-0.01]), torch.tensor([0.04])
    gc.action_space = SimpleNamespace(shape=(B, 1))
    ga = (torch.rand((B, 1), generator=g) * 3 - 1.5).float()
    gc.set_action(ga.clone())
    out.update(gripper_action=ga.numpy(), gripper_target_qpos=gc.articulation.written[0].numpy())
    # ---------------- Fetch arm (11 ancestors, 7 controlled): reference compute_ik on the masked J ----
    # The stock reference returns an un-scatterable (B, 11) result here; the golden
    # vector is the reference expression q0 + pinv(J) @ a evaluated by the reference's code path on
    # the intended masked Jacobian (J[:, :, controlled]).
    fframes = O.load_chain(os.path.join(assets, "fetch_kin.urdf"), "gripper_link")
    lim = np.clip(O.joint_limits(fframes), -3, 3)
    qf = torch.from_numpy(np.random.default_rng(7).uniform(lim[:, 0] * 0.8, lim[:, 1] * 0.8, (B, 11))).float()
    ctrl = list(range(4, 11))
    kf = object.__new__(kinmod.Kinematics)
    kf.use_gpu_ik, kf.device, kf.active_ancestor_joint_idxs = True, qf.device, ctrl

    def fjac(q7):
        full = qf.double().numpy().copy()
        full[:, ctrl] = q7.double().numpy()
        J, _, _ = O.jacobian(fframes, full)
        return torch.from_numpy(J[:, :, ctrl]).to(q7.dtype)

    kf.pk_chain = SimpleNamespace(jacobian=fjac)
    fa = a_scaled.clone()
    out.update(fetch_q=qf.numpy(), fetch_action=fa.numpy(),
               fetch_target_qpos=kf.compute_ik(None, qf, action=fa, use_delta_ik_solver=True).numpy())
    np.savez_compressed(OUT, **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", os.path.relpath(OUT), os.path.getsize(OUT), "B;", {k: np.asarray(v).shape for k, v in out.items()})
    return 0


if __name__ == "__main__":
    sys.exit(main())
