"""Stand-in for the SAPIEN / PhysX objects the EE controllers touch.

SAPIEN needs a Vulkan loader even for physics-only use (reference
mani_skill/envs/sapien_env.py:1086-1088) and is not installable here, so tests,
``bench.py`` and ``smoke()`` drive the controllers through this shim.  It exposes
exactly the attributes listed in SURVEY.md section 8(b)-3 with the PhysX GPU buffer
layouts:

* ``px.cuda_articulation_qpos / _qvel / _target_qpos / _target_qvel``  (M, max_dof) fp32
  (reference utils/structs/articulation.py:613-622, 731-738)
* ``px.cuda_rigid_body_data``  (Nbody, 13) fp32, [:7] = p xyz, q wxyz
  (reference utils/structs/link.py:225-237)

each wrapped so that ``.torch()`` returns the backing CUDA tensor.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from maniskill_b200.structs.pose import Pose
from maniskill_b200.urdf import UrdfJoint, parse_urdf


class CudaArray:
    """``sapien`` CudaArray look-alike: ``.torch()`` gives the tensor view."""

    def __init__(self, t: torch.Tensor):
        self._t = t

    def torch(self) -> torch.Tensor:
        return self._t


class FakePhysxSystem:
    def __init__(self, n_art: int, max_dof: int, n_bodies: int, device, timestep: float = 1.0 / 100):
        self.timestep = timestep
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=device)
        self.cuda_articulation_qpos = CudaArray(z(n_art, max_dof))
        self.cuda_articulation_qvel = CudaArray(z(n_art, max_dof))
        self.cuda_articulation_target_qpos = CudaArray(z(n_art, max_dof))
        self.cuda_articulation_target_qvel = CudaArray(z(n_art, max_dof))
        body = z(n_bodies, 13)
        body[:, 3] = 1.0
        self.cuda_rigid_body_data = CudaArray(body)
        self.applied_target_position = 0

    def gpu_apply_articulation_target_position(self):
        self.applied_target_position += 1


class FakeJoint:
    def __init__(self, name: str, jtype: str, limits, active_index: Optional[int], device):
        self.name = name
        self.type = jtype
        self.limits = limits
        self.active_index = None if active_index is None else torch.tensor([active_index], dtype=torch.int32, device=device)
        self.parent_link: Optional["FakeLink"] = None
        self.child_link: Optional["FakeLink"] = None
        self.drive_properties = None
        self.friction = None

    def set_drive_properties(self, stiffness, damping, force_limit=None, mode="force"):
        self.drive_properties = (float(stiffness), float(damping), None if force_limit is None else float(force_limit), mode)

    def set_friction(self, friction):
        self.friction = float(friction)


class FakeLink:
    def __init__(self, name: str, articulation: "FakeArticulation"):
        self.name = name
        self.articulation = articulation
        self.joint: Optional[FakeJoint] = None
        self._body_data_index: Optional[torch.Tensor] = None  # (num_envs,) int64 rows of rigid body data (structs/base.py:104-110)

    @property
    def pose(self) -> Pose:
        # utils/structs/link.py:225-237 (scene offsets are zero in the shim)
        raw = self.articulation.px.cuda_rigid_body_data.torch()[self._body_data_index.long(), :7]
        return Pose.create(raw)


class FakeScene:
    def __init__(self, num_envs: int, device):
        self.num_envs = num_envs
        self.gpu_sim_enabled = True
        self.parallel_in_single_scene = False
        self.device = device
        self._reset_mask = torch.ones(num_envs, dtype=torch.bool, device=device)


class FakeArticulation:
    """One robot per sub-scene, ``num_envs`` copies, PhysX row layout.

    ``active_joint_order``: names of the active (movable) joints in PhysX dof order;
    default = URDF file order of the movable joints.
    ``n_pad_rows``: extra articulation rows placed BEFORE ours in the PhysX buffers (another
    articulation in the scene) so that ``_data_index`` is not the identity.
    ``max_dof``: row stride of the PhysX buffers (>= own dof)."""

    def __init__(self, urdf_path: str, num_envs: int, device="cuda", active_joint_order: Optional[Sequence[str]] = None,
                 max_dof: Optional[int] = None, n_pad_rows: int = 0, interleave_rows: bool = False,
                 ee_link_names: Sequence[str] = (), name: str = "robot"):
        self.name = name
        self.device = torch.device(device)
        self.urdf_path = urdf_path
        link_names, joints = parse_urdf(urdf_path)
        self._urdf_joints: Dict[str, UrdfJoint] = {j.name: j for j in joints}
        movable = [j for j in joints if j.movable]
        if active_joint_order is None:
            active_joint_order = [j.name for j in movable]
        assert sorted(active_joint_order) == sorted(j.name for j in movable)
        self._active_names = list(active_joint_order)
        self.dof = len(self._active_names)
        self.max_dof = max_dof or self.dof
        assert self.max_dof >= self.dof
        self.scene = FakeScene(num_envs, self.device)
        self.num_envs = num_envs

        # links / joints graph
        self.links: List[FakeLink] = [FakeLink(n, self) for n in link_names]
        self.links_map: Dict[str, FakeLink] = {l.name: l for l in self.links}
        children = {j.child for j in joints}
        roots = [l for l in self.links if l.name not in children]
        assert len(roots) == 1, [r.name for r in roots]
        self.root = roots[0]
        root_joint = FakeJoint("", "fixed", None, None, self.device)  # SAPIEN gives the root a dummy joint
        root_joint.child_link = self.root
        self.root.joint = root_joint
        self.joints: List[FakeJoint] = [root_joint]
        self.joints_map: Dict[str, FakeJoint] = {}
        for j in joints:
            ai = self._active_names.index(j.name) if j.name in self._active_names else None
            fj = FakeJoint(j.name, j.type, (j.lower, j.upper), ai, self.device)
            fj.parent_link = self.links_map[j.parent]
            fj.child_link = self.links_map[j.child]
            fj.child_link.joint = fj
            self.joints.append(fj)
            self.joints_map[j.name] = fj
        self.active_joints = [self.joints_map[n] for n in self._active_names]

        # PhysX buffers
        if interleave_rows:
            n_art = 2 * num_envs
            rows = torch.arange(num_envs, device=self.device, dtype=torch.int32) * 2 + 1
        else:
            n_art = n_pad_rows + num_envs
            rows = torch.arange(num_envs, device=self.device, dtype=torch.int32) + n_pad_rows
        self._data_index = rows
        self._identity_rows = bool(n_pad_rows == 0 and not interleave_rows)
        tracked = [self.root.name] + [n for n in ee_link_names if n != self.root.name]
        self.px = FakePhysxSystem(n_art, self.max_dof, len(tracked) * num_envs, self.device)
        for k, lname in enumerate(tracked):
            self.links_map[lname]._body_data_index = torch.arange(num_envs, device=self.device, dtype=torch.int64) + k * num_envs
        self._cached_joint_target_indices = dict()

        lim = np.zeros((self.max_dof, 2), dtype=np.float32)
        for i, n in enumerate(self._active_names):
            j = self._urdf_joints[n]
            lim[i] = (max(j.lower, -1e30), min(j.upper, 1e30))
        self.qlimits = torch.from_numpy(lim).to(self.device)[None].repeat(num_envs, 1, 1)

    # -- reference Articulation API used by the controllers --------------------------------
    def get_active_joints(self):
        return self.active_joints

    def get_qpos(self) -> torch.Tensor:
        return self.qpos

    @property
    def qpos(self) -> torch.Tensor:
        # articulation.py:613-622: advanced-index gather (a copy)
        return self.px.cuda_articulation_qpos.torch()[self._data_index.long(), : self.max_dof]

    @qpos.setter
    def qpos(self, v: torch.Tensor):
        self.px.cuda_articulation_qpos.torch()[self._data_index.long(), : v.shape[-1]] = v.to(self.device).float()

    def get_qvel(self) -> torch.Tensor:
        return self.px.cuda_articulation_qvel.torch()[self._data_index.long(), : self.max_dof]

    def get_qlimits(self) -> torch.Tensor:
        return self.qlimits

    @property
    def pose(self) -> Pose:
        return self.root.pose

    def set_root_pose(self, pose: Pose):
        self.px.cuda_rigid_body_data.torch()[self.root._body_data_index.long(), :7] = pose.raw_pose

    def set_link_pose(self, link_name: str, pose: Pose):
        l = self.links_map[link_name]
        self.px.cuda_rigid_body_data.torch()[l._body_data_index.long(), :7] = pose.raw_pose

    def set_joint_drive_targets(self, targets, joints=None, joint_indices: torch.Tensor = None):
        # articulation.py:720-738 (GPU branch)
        targets = torch.as_tensor(targets, device=self.device)
        key = joint_indices
        if key not in self._cached_joint_target_indices:
            self._cached_joint_target_indices[key] = torch.meshgrid(self._data_index.long(), joint_indices.long(), indexing="ij")
        gx, gy = self._cached_joint_target_indices[key]
        self.px.cuda_articulation_target_qpos.torch()[gx, gy] = targets

    def set_joint_drive_velocity_targets(self, targets, joints=None, joint_indices: torch.Tensor = None):
        # articulation.py:743-761 (GPU branch)
        targets = torch.as_tensor(targets, device=self.device)
        key = joint_indices
        if key not in self._cached_joint_target_indices:
            self._cached_joint_target_indices[key] = torch.meshgrid(self._data_index.long(), joint_indices.long(), indexing="ij")
        gx, gy = self._cached_joint_target_indices[key]
        self.px.cuda_articulation_target_qvel.torch()[gx, gy] = targets

    @property
    def drive_velocity_targets(self) -> torch.Tensor:
        return self.px.cuda_articulation_target_qvel.torch()[self._data_index.long(), : self.max_dof]

    @property
    def drive_targets(self) -> torch.Tensor:
        return self.px.cuda_articulation_target_qpos.torch()[self._data_index.long(), : self.max_dof]
