"""Batched Pose, the I/O type of ``Kinematics.compute_ik``.

API mirror of mani_skill/utils/structs/pose.py:30-273 restricted to what the
controller path touches: raw (N, 7) = [p xyz, q wxyz], ``create`` / ``create_from_pq``,
``p`` / ``q`` views, ``__mul__``, ``inv``, ``to_transformation_matrix``, indexing.
(sapien.Pose inputs are not supported: SAPIEN is not a dependency of this package.)
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence, Union

import numpy as np
import torch

from maniskill_b200 import geometry as G

Array = Union[torch.Tensor, np.ndarray, Sequence]


def _to_tensor(x, device=None) -> torch.Tensor:
    # common.to_tensor semantics (mani_skill/utils/common.py:146-174): float64 -> float32
    if isinstance(x, torch.Tensor):
        t = x if device is None else x.to(device)
    else:
        t = torch.as_tensor(np.asarray(x), device=device)
    if t.dtype == torch.float64:
        t = t.float()
    return t


def _batched(x, device=None) -> Optional[torch.Tensor]:
    if x is None:
        return None
    t = _to_tensor(x, device)
    return t[None, :] if t.dim() == 1 else t


@dataclass
class Pose:
    raw_pose: torch.Tensor

    @classmethod
    def create_from_pq(cls, p: Optional[Array] = None, q: Optional[Array] = None, device=None) -> "Pose":
        if device is None:
            device = p.device if isinstance(p, torch.Tensor) else q.device if isinstance(q, torch.Tensor) else None
        if p is None:
            p = torch.zeros((1, 3), device=device)
        if q is None:
            q = torch.zeros((1, 4), device=device)
            q[:, 0] = 1
        p, q = _batched(p, device), _batched(q, device)
        if p.shape[0] > q.shape[0]:
            assert q.shape[0] == 1
            q = q.repeat(p.shape[0], 1)
        elif p.shape[0] < q.shape[0]:
            assert p.shape[0] == 1
            p = p.repeat(q.shape[0], 1)
        return cls(raw_pose=torch.hstack([p, q]))

    @classmethod
    def create(cls, pose: Union["Pose", Array], device=None) -> "Pose":
        if isinstance(pose, cls):
            return cls(raw_pose=pose.raw_pose if device is None else pose.raw_pose.to(device))
        t = _batched(pose, device)
        assert t.dim() == 2, t.shape
        if t.shape[-1] == 3:
            return cls.create_from_pq(p=t, device=t.device)
        assert t.shape[-1] == 7, t.shape
        return cls(raw_pose=t)

    def __getitem__(self, i) -> "Pose":
        return Pose.create(self.raw_pose[i])

    def __len__(self) -> int:
        return len(self.raw_pose)

    @property
    def shape(self):
        return self.raw_pose.shape

    @property
    def device(self):
        return self.raw_pose.device

    def to(self, device) -> "Pose":
        if self.raw_pose.device == torch.device(device):
            return self
        return Pose.create(self.raw_pose.to(device))

    def __mul__(self, other: Union["Pose", Array]) -> "Pose":
        other = Pose.create(other, device=self.device)
        a = self
        if len(other) == 1 and len(a) > 1:
            other = Pose(other.raw_pose.repeat(len(a), 1))
        elif len(a) == 1 and len(other) > 1:
            a = Pose(a.raw_pose.repeat(len(other), 1))
        q = G.quat_mul(a.q, other.q)
        p = a.p + G.quat_apply(a.q, other.p)
        return Pose.create_from_pq(p, q)

    def inv(self) -> "Pose":
        qi = G.quat_conj(self.q)
        return Pose(torch.cat((G.quat_apply(qi, -self.p), qi), -1))

    def to_transformation_matrix(self) -> torch.Tensor:
        b = self.raw_pose.shape[0]
        m = torch.zeros((b, 4, 4), device=self.device, dtype=self.raw_pose.dtype)
        m[:, :3, :3] = G.quat_to_matrix(self.q)
        m[:, :3, 3] = self.p
        m[:, 3, 3] = 1
        return m

    def get_p(self):
        return self.p

    def get_q(self):
        return self.q

    def set_p(self, p):
        self.p = p

    def set_q(self, q):
        self.q = q

    @property
    def p(self) -> torch.Tensor:
        return self.raw_pose[..., :3]

    @p.setter
    def p(self, v):
        self.raw_pose[..., :3] = _to_tensor(v, self.device)

    @property
    def q(self) -> torch.Tensor:
        return self.raw_pose[..., 3:]

    @q.setter
    def q(self, v):
        self.raw_pose[..., 3:] = _to_tensor(v, self.device)


def vectorize_pose(pose: Union[Pose, Array], device=None) -> torch.Tensor:
    if isinstance(pose, Pose):
        return pose.raw_pose if device is None else pose.raw_pose.to(device)
    return _to_tensor(pose, device)
