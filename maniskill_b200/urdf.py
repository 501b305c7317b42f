"""URDF -> flattened joint table ("chain flattening", done once per controller).

The reference builds its GPU chain with
``pk.build_serial_chain_from_urdf(urdf_str, end_link_name=...)``
(mani_skill/agents/controllers/utils/kinematics.py:106-109) and discovers the
active ancestor joints by walking ``link.joint.parent_link`` from the end link
to the root (kinematics.py:54-61).  Both reduce to: the serial chain of URDF
joints from the root link to ``end_link_name``.  This module parses the URDF
with the stdlib XML parser and folds that chain into the canonical form the
CUDA kernels consume:

    T_ee(q) = G_1 Z_1(q_1) G_2 Z_2(q_2) ... G_n Z_n(q_n) G_tool

where every movable joint acts about/along the local **z** axis
(``Z_i`` = Rz(q_i) for revolute/continuous, Tz(q_i) for prismatic) and every
``G_i = (M_i, t_i)`` is a constant rigid transform that absorbs the URDF joint
origin, all fixed joints in between and the change of basis that maps the URDF
joint axis onto z.  For Panda / Fetch every ``M_i`` is a signed permutation,
which the kernels constant-fold.
"""
from __future__ import annotations

import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

JOINT_FIXED = 0
JOINT_REVOLUTE = 1  # revolute and continuous
JOINT_PRISMATIC = 2

MSIK_MAX_JOINTS = 16  # must match include/msik.h


@dataclass
class UrdfJoint:
    name: str
    type: str
    parent: str
    child: str
    xyz: np.ndarray
    rpy: np.ndarray
    axis: np.ndarray
    lower: float
    upper: float
    mimic: Optional[str] = None

    @property
    def movable(self) -> bool:
        return self.type in ("revolute", "continuous", "prismatic")


def _floats(s: Optional[str], n: int, default: Sequence[float]) -> np.ndarray:
    if s is None:
        return np.asarray(default, dtype=np.float64)
    v = np.asarray([float(x) for x in s.split()], dtype=np.float64)
    if v.shape != (n,):
        raise ValueError(f"expected {n} floats, got {s!r}")
    return v


def parse_urdf(urdf: "str | bytes") -> Tuple[List[str], List[UrdfJoint]]:
    """Parse URDF text (or a path) into (link names, joints)."""
    if isinstance(urdf, bytes):
        text = urdf.decode("utf-8")
    else:
        text = urdf
    if "<" not in text:  # a path
        with open(text, "r", encoding="utf-8") as f:
            text = f.read()
    root = ET.fromstring(text)
    links = [l.get("name") for l in root.findall("link")]
    joints: List[UrdfJoint] = []
    for j in root.findall("joint"):
        origin = j.find("origin")
        axis = j.find("axis")
        limit = j.find("limit")
        mimic = j.find("mimic")
        jtype = j.get("type")
        lower, upper = -math.inf, math.inf
        if limit is not None and jtype != "continuous":
            lower = float(limit.get("lower", "-inf"))
            upper = float(limit.get("upper", "inf"))
        joints.append(
            UrdfJoint(
                name=j.get("name"),
                type=jtype,
                parent=j.find("parent").get("link"),
                child=j.find("child").get("link"),
                xyz=_floats(origin.get("xyz") if origin is not None else None, 3, (0, 0, 0)),
                rpy=_floats(origin.get("rpy") if origin is not None else None, 3, (0, 0, 0)),
                # URDF default axis is (1, 0, 0)
                axis=_floats(axis.get("xyz") if axis is not None else None, 3, (1, 0, 0)),
                lower=lower,
                upper=upper,
                mimic=mimic.get("joint") if mimic is not None else None,
            )
        )
    return links, joints


def serial_chain(joints: Sequence[UrdfJoint], end_link: str) -> List[UrdfJoint]:
    """URDF joints from the root link to ``end_link`` in root->end order
    (same walk as kinematics.py:54-61, on URDF data instead of SAPIEN links)."""
    by_child: Dict[str, UrdfJoint] = {j.child: j for j in joints}
    known_links = set(by_child) | {j.parent for j in joints}
    if end_link not in known_links:
        raise KeyError(f"link {end_link!r} not found in URDF")
    out: List[UrdfJoint] = []
    cur = end_link
    while cur in by_child:
        j = by_child[cur]
        out.append(j)
        cur = j.parent
        if len(out) > 1024:
            raise ValueError("URDF joint graph has a cycle")
    return out[::-1]


def rpy_matrix(rpy: np.ndarray) -> np.ndarray:
    """URDF fixed-axis roll/pitch/yaw: R = Rz(yaw) Ry(pitch) Rx(roll)."""
    r, p, y = (float(v) for v in rpy)
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array(
        [
            [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
            [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
            [-sp, cp * sr, cp * cr],
        ],
        dtype=np.float64,
    )


def _z_to_axis(a: np.ndarray) -> np.ndarray:
    """A rotation A with A @ z == a.  Axis-aligned inputs give signed permutations."""
    a = a / np.linalg.norm(a)
    z = np.array([0.0, 0.0, 1.0])
    c = float(a @ z)
    if c > 1.0 - 1e-12:
        return np.eye(3)
    if c < -1.0 + 1e-12:
        return np.diag([1.0, -1.0, -1.0])  # Rx(pi)
    # signed-permutation choices for +-x / +-y keep the kernel tables sparse
    for k in (0, 1):
        if abs(abs(a[k]) - 1.0) < 1e-12:
            s = math.copysign(1.0, a[k])
            A = np.zeros((3, 3))
            A[k, 2] = s  # z -> s * e_k
            o = 1 - k
            # complete to a proper rotation: columns (x', y', z')
            if k == 0:
                # z -> s*x ; choose y' = y, x' = y' cross z'
                A[1, 1] = 1.0
                A[:, 0] = np.cross(A[:, 1], A[:, 2])
            else:
                # z -> s*y ; choose x' = x, y' = z' cross x'
                A[0, 0] = 1.0
                A[:, 1] = np.cross(A[:, 2], A[:, 0])
            return A
    v = np.cross(z, a)
    s2 = float(v @ v)
    vx = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])
    return np.eye(3) + vx + vx @ vx * ((1.0 - c) / s2)


def _snap(m: np.ndarray, tol: float = 1e-12) -> np.ndarray:
    """Snap |x|<tol to 0 and |x|-1<tol to +-1 (cos(pi/2) = 6e-17 in URDF rpy)."""
    m = np.where(np.abs(m) < tol, 0.0, m)
    m = np.where(np.abs(np.abs(m) - 1.0) < tol, np.sign(m), m)
    return m


@dataclass
class ChainTable:
    """Flattened chain: canonical z-axis joints with constant pre-transforms."""

    joint_names: List[str]
    joint_types: np.ndarray  # (n,) int32 in {JOINT_REVOLUTE, JOINT_PRISMATIC}
    M: np.ndarray  # (n, 3, 3) float64
    t: np.ndarray  # (n, 3) float64
    M_tool: np.ndarray  # (3, 3)
    t_tool: np.ndarray  # (3,)
    lower: np.ndarray  # (n,)
    upper: np.ndarray  # (n,)
    end_link: str = ""
    root_link: str = ""
    raw_frames: List[UrdfJoint] = field(default_factory=list)

    @property
    def n(self) -> int:
        return len(self.joint_names)

    def packed_f32(self) -> np.ndarray:
        """Row-major float32 blob in the layout of ``msik_chain_desc`` (include/msik.h):
        per joint 12 floats (M row-major 9, t 3), then tool 12 floats."""
        out = np.zeros((self.n + 1, 12), dtype=np.float32)
        for i in range(self.n):
            out[i, :9] = self.M[i].reshape(9)
            out[i, 9:] = self.t[i]
        out[self.n, :9] = self.M_tool.reshape(9)
        out[self.n, 9:] = self.t_tool
        return out


def flatten_chain(frames: Sequence[UrdfJoint], end_link: str = "") -> ChainTable:
    """Fold fixed joints / origins / axis changes into per-joint constants."""
    names: List[str] = []
    types: List[int] = []
    Ms: List[np.ndarray] = []
    ts: List[np.ndarray] = []
    lo: List[float] = []
    hi: List[float] = []
    # pending fixed transform accumulated since the previous movable joint
    R_acc = np.eye(3)
    p_acc = np.zeros(3)
    for j in frames:
        R_off = rpy_matrix(j.rpy)
        # acc <- acc * (R_off, xyz)
        p_acc = p_acc + R_acc @ j.xyz
        R_acc = R_acc @ R_off
        if not j.movable:
            if j.type != "fixed":
                raise NotImplementedError(f"joint type {j.type!r} ({j.name}) is not supported")
            continue
        if np.linalg.norm(j.axis) < 1e-12:
            raise ValueError(f"joint {j.name} has a zero axis")
        A = _z_to_axis(j.axis)
        names.append(j.name)
        types.append(JOINT_PRISMATIC if j.type == "prismatic" else JOINT_REVOLUTE)
        Ms.append(_snap(R_acc @ A))
        ts.append(_snap(p_acc.copy()))
        lo.append(j.lower)
        hi.append(j.upper)
        # after Z(q): undo the change of basis
        R_acc = A.T.copy()
        p_acc = np.zeros(3)
    n = len(names)
    if n == 0:
        raise ValueError("chain has no movable joints")
    if n > MSIK_MAX_JOINTS:
        raise ValueError(f"chain has {n} movable joints; MSIK_MAX_JOINTS={MSIK_MAX_JOINTS}")
    return ChainTable(
        joint_names=names,
        joint_types=np.asarray(types, dtype=np.int32),
        M=np.stack(Ms),
        t=np.stack(ts),
        M_tool=_snap(R_acc),
        t_tool=_snap(p_acc),
        lower=np.asarray(lo, dtype=np.float64),
        upper=np.asarray(hi, dtype=np.float64),
        end_link=end_link,
        root_link=frames[0].parent if frames else "",
        raw_frames=list(frames),
    )


def build_chain_table(urdf: "str | bytes", end_link_name: str) -> ChainTable:
    """Counterpart of ``pk.build_serial_chain_from_urdf`` (kinematics.py:106-109)."""
    _, joints = parse_urdf(urdf)
    frames = serial_chain(joints, end_link_name)
    return flatten_chain(frames, end_link=end_link_name)
