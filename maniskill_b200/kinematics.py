"""Kinematics: drop-in for mani_skill/agents/controllers/utils/kinematics.py.

Same constructor and ``compute_ik`` signature (kinematics.py:31-37,124-141), but the
arithmetic that the reference delegates to pytorch_kinematics (GPU) runs in the
hand-written sm_100a kernels behind libmsik.so (include/msik.h).  There is no
CPU path: constructing this on a non-CUDA articulation raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import numpy as np
import torch

from maniskill_b200 import _lib as L
from maniskill_b200.structs.pose import Pose
from maniskill_b200.urdf import ChainTable, build_chain_table


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class Kinematics:
    """Kinematics solver bound to one articulation view (one robot per sub-scene).

    Attributes kept from the reference: ``urdf_path``, ``end_link``, ``end_link_idx``,
    ``active_joint_indices``, ``articulation``, ``device``, ``active_ancestor_joints``,
    ``active_ancestor_joint_idxs``, ``controlled_joints_idx_in_qmask``, ``use_gpu_ik``,
    ``qmask``.
    """

    # recalled pytorch_kinematics defaults (UNVERIFIED); tree values:
    # max_iterations=200, num_retries=1, early_stopping_any_converged=True (kinematics.py:110-117)
    max_iterations: int = 200
    lr: float = 0.2
    pos_tolerance: float = 1e-3
    rot_tolerance: float = 1e-2
    no_improvement_patience: int = 2
    # pk `regularlization` (sic): lambda^2 of J^T (J J^T + lambda^2 I)^-1 dx.  0 selects the undamped limit everywhere;
    # with a positive value envs near a singularity take the exact damped update (include/msik.h msik_iter_cfg.damping)
    regularization: float = 1e-9
    guard_pivot: float = 0.0  # 0: derived from `regularization`; 1e30: exact damped update in every iteration
    ik_mapping: str = "auto"  # "auto" | "thread" (one thread per env) | "lanes" (8 lanes per env, chains of <= 8 joints)

    def __init__(self, urdf_path: str, end_link_name: str, articulation, active_joint_indices: torch.Tensor,
                 specialise: Optional[bool] = None):
        """Reference signature (kinematics.py:31-37) plus ``specialise``: compile straight-line FK / Jacobian
        kernels for this URDF chain at construction when the library has none built in (maniskill_b200/jit.py;
        default: the MSIK_JIT=1 environment variable; otherwise such chains run the table-driven kernels)."""
        self.specialise = (os.environ.get("MSIK_JIT", "0") == "1") if specialise is None else bool(specialise)
        self.urdf_path = urdf_path
        self.end_link = articulation.links_map[end_link_name]
        self.end_link_idx = articulation.links.index(self.end_link)
        self.active_joint_indices = active_joint_indices
        self.articulation = articulation
        self.device = articulation.device
        # ancestor walk, kinematics.py:54-61
        cur_link = self.end_link
        active_ancestor_joints = []
        while cur_link is not None:
            if cur_link.joint.active_index is not None:
                active_ancestor_joints.append(cur_link.joint)
            cur_link = cur_link.joint.parent_link
        self.active_ancestor_joints = active_ancestor_joints[::-1]
        # kinematics.py:65-71
        self.active_ancestor_joint_idxs: List[int] = [int(x.active_index[0]) for x in self.active_ancestor_joints]
        self.controlled_joints_idx_in_qmask = [
            self.active_ancestor_joint_idxs.index(int(idx)) for idx in self.active_joint_indices
        ]
        if torch.device(self.device).type != "cuda":
            raise L.MsikError(
                "maniskill_b200.Kinematics is the B200 CUDA path of the controller; "
                f"articulation.device={self.device!r} has no implementation here (no CPU fallback)."
            )
        self.use_gpu_ik = True
        self._setup_gpu()

    # ------------------------------------------------------------------ setup
    def _setup_gpu(self):
        """kinematics.py:91-122: flatten the URDF chain, upload it, build qmask."""
        self._lib = L.load()
        L.require_device()
        with open(self.urdf_path, "rb") as f:
            urdf_str = f.read()
        self.chain_table: ChainTable = build_chain_table(urdf_str, self.end_link.name)
        n = self.chain_table.n
        if n != len(self.active_ancestor_joints):
            raise L.MsikError(
                f"URDF chain to {self.end_link.name!r} has {n} movable joints but the articulation "
                f"reports {len(self.active_ancestor_joints)} active ancestors"
            )
        names = [j.name for j in self.active_ancestor_joints]
        if names != self.chain_table.joint_names:
            raise L.MsikError(f"ancestor joints {names} do not match URDF chain {self.chain_table.joint_names}")
        ctrl = sorted(self.controlled_joints_idx_in_qmask)
        if ctrl != list(self.controlled_joints_idx_in_qmask):
            # the reference scatters column j of the result to active_joint_indices[j]
            # (pd_joint_pos.py:68-71), i.e. it silently assumes chain order
            raise L.MsikError("controlled joints must be listed in chain order (root -> end link)")
        if self.specialise:
            from maniskill_b200 import jit

            mask = 0
            for i in self.controlled_joints_idx_in_qmask:
                mask |= 1 << i
            jit.specialise(self.urdf_path, self.end_link.name, mask)  # built-in chains still win in msik_chain_create
        desc = L.ChainDesc()
        desc.n_joints = n
        packed = self.chain_table.packed_f32()
        for i in range(n):
            desc.joint_type[i] = int(self.chain_table.joint_types[i])
            desc.lower[i] = float(np.float32(self.chain_table.lower[i]))
            desc.upper[i] = float(np.float32(self.chain_table.upper[i]))
        for i in range(n + 1):
            for k in range(12):
                desc.pre[i][k] = float(packed[i, k])
        self._desc = desc
        handle = C.c_void_p()
        L.check(self._lib.msik_chain_create(C.byref(desc), C.byref(handle)))
        self._chain = handle
        buf = C.create_string_buffer(64)
        L.check(self._lib.msik_chain_kernel_name(self._chain, buf, 64))
        self.kernel_name = buf.value.decode()
        self.ctrl_mask = 0
        for i in self.controlled_joints_idx_in_qmask:
            self.ctrl_mask |= 1 << i
        self.n_ctrl = len(self.controlled_joints_idx_in_qmask)
        self.qmask = torch.zeros(len(self.active_ancestor_joints), dtype=torch.bool, device=self.device)
        self.qmask[self.controlled_joints_idx_in_qmask] = 1
        self.joint_limits = torch.tensor(
            np.stack([self.chain_table.lower, self.chain_table.upper], 1), dtype=torch.float32, device=self.device
        )

    def __del__(self):
        try:
            for h in getattr(self, "_pipes", {}).values():
                self._lib.msik_pipeline_destroy(h)
            self._pipes = {}
            if getattr(self, "_chain", None):
                self._lib.msik_chain_destroy(self._chain)
                self._chain = None
        except Exception:
            pass

    # ------------------------------------------------------------------ cfg helpers
    def step_cfg(self, n_envs: int, qpos_stride: int, target_stride: int, action_stride: int,
                 action_offset: int, flags: int, pos_low=(0, 0, 0), pos_high=(0, 0, 0), rot_scale: float = 1.0,
                 col_shift: int = 0) -> L.StepCfg:
        cfg = L.StepCfg()
        cfg.n_envs = n_envs
        cfg.qpos_stride = qpos_stride
        cfg.target_stride = target_stride
        cfg.action_stride = action_stride
        cfg.action_offset = action_offset
        cfg.flags = flags
        cfg.ctrl_mask = self.ctrl_mask
        for i, c in enumerate(self.active_ancestor_joint_idxs):
            cfg.anc_col[i] = c + col_shift
        for k in range(3):
            cfg.pos_low[k] = float(pos_low[k])
            cfg.pos_high[k] = float(pos_high[k])
        cfg.rot_scale = float(rot_scale)
        return cfg

    def iter_cfg(self) -> L.IterCfg:
        it = L.IterCfg()
        it.max_iterations = int(self.max_iterations)
        it.patience = -1 if self.no_improvement_patience is None else int(self.no_improvement_patience)
        it.lr = float(self.lr)
        it.pos_tol = float(self.pos_tolerance)
        it.rot_tol = float(self.rot_tolerance)
        it.rank_rtol = 0.0
        it.damping = float(self.regularization)
        it.guard_pivot = float(self.guard_pivot)
        it.mapping = {"auto": L.IK_AUTO, "thread": L.IK_THREAD_PER_ENV, "lanes": L.IK_LANE_PER_JOINT}[self.ik_mapping]
        return it

    @staticmethod
    def _rows(t: torch.Tensor) -> torch.Tensor:
        """fp32 CUDA 2-D tensor with unit inner stride (row stride may be padded)."""
        assert t.is_cuda and t.dim() == 2, (t.device, t.shape)
        if t.dtype != torch.float32:
            t = t.float()
        if t.stride(1) != 1 or (t.shape[0] > 1 and t.stride(0) < t.shape[1]):
            t = t.contiguous()
        return t

    @staticmethod
    def _row_stride(t: torch.Tensor) -> int:
        return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])

    @staticmethod
    def _base_view(t: torch.Tensor):
        """(base pointer, row stride, first column) of a 2-D tensor that may be a column slice
        of a wider row-major buffer.  The kernels move whole rows of `row stride` floats, so
        they are handed the start of the parent row, never a pointer into the middle of it."""
        stride = Kinematics._row_stride(t)
        col = 0
        if stride > t.shape[1]:
            col = t.storage_offset() % stride
            if col + t.shape[1] > stride:  # not a column slice of full rows: cannot rebase
                t = t.contiguous()
                return t, t.data_ptr(), t.shape[1], 0
        return t, t.data_ptr() - 4 * col, stride, col

    # ------------------------------------------------------------------ API
    def compute_ik(self, target_pose: Pose, q0: torch.Tensor, pos_only: bool = False, action=None,
                   use_delta_ik_solver: bool = False, *, return_info: bool = False):
        """Same contract as kinematics.py:124-186 (GPU branch).

        ``return_info`` (iterative path only, not in the reference): also return ``{"iterations": (B,) int32 -- negative
        where the env finished on the exact damped step --, "error": (B, 2) final position / rotation error norms}``.

        Returns the joint targets of the controlled joints, shape (B, c).  For robots whose
        controlled joints are all active ancestors (Panda) that is exactly the reference's
        (B, n) result; for chains with uncontrolled ancestors (Fetch) the reference returns
        an un-scatterable (B, n) tensor and we return the masked (B, c) one.
        """
        q0 = self._rows(q0)
        B = q0.shape[0]
        out = torch.empty((B, self.n_ctrl), dtype=torch.float32, device=q0.device)
        q0, qptr, qs, qcol = self._base_view(q0)
        if use_delta_ik_solver:
            assert action is not None, "delta IK needs the (pre-processed) action"
            action = self._rows(action)
            nd = 3 if pos_only else 6
            assert action.shape == (B, nd), (action.shape, (B, nd))
            action, aptr, astride, acol = self._base_view(action)
            flags = L.F_USE_DELTA | L.F_NO_DRIVE | (L.F_POS_ONLY if pos_only else 0)
            cfg = self.step_cfg(B, qs, qs, astride, acol, flags, col_shift=qcol)
            L.check(self._lib.msik_delta_step(self._chain, C.byref(cfg), qptr, None, aptr, None, L.ptr(out), None, _stream()))
            return out
        tp = self._rows(Pose.create(target_pose, device=q0.device).raw_pose)
        if tp.shape[0] == 1 and B > 1:
            tp = tp.repeat(B, 1)
        tp = tp.contiguous()
        assert tp.shape == (B, 7), tp.shape
        flags = L.F_TARGET_GIVEN | L.F_NO_DRIVE
        cfg = self.step_cfg(B, qs, qs, 0, 0, flags, col_shift=qcol)
        it = self.iter_cfg()
        iters = torch.empty(B, dtype=torch.int32, device=q0.device) if return_info else None
        err = torch.empty((B, 2), dtype=torch.float32, device=q0.device) if return_info else None
        L.check(self._lib.msik_target_ik(self._chain, C.byref(cfg), C.byref(it), qptr, None, None, L.ptr(tp),
                                         None, None, L.ptr(out), None, L.ptr(iters), L.ptr(err), _stream()))
        if return_info:
            return out, {"iterations": iters, "error": err}
        return out

    def forward_kinematics(self, q0: torch.Tensor, return_jacobian: bool = False, root_pose=None):
        """EE pose in the chain root frame (and optionally the (B, 6, n) base-frame Jacobian,
        rows [linear; angular]) -- what pk ``chain.jacobian(q, ret_eef_pose=True)`` returns.
        With ``root_pose`` (Pose or (B,7)) the result is the WORLD pose ``root_pose * FK(q0)`` from the same
        launch (``msik_fk_world``): the EE / tcp pose for proprioception without a PhysX read-back
        (agents/robots/panda/panda.py:232-234) and the FK of trajectory/utils/actions/conversion.py:136-143."""
        q0 = self._rows(q0)
        B = q0.shape[0]
        n = self.chain_table.n
        pose = torch.empty((B, 7), dtype=torch.float32, device=q0.device)
        jac = torch.empty((B, 6, n), dtype=torch.float32, device=q0.device) if return_jacobian else None
        q0, qptr, qs, qcol = self._base_view(q0)
        cfg = self.step_cfg(B, qs, qs, 0, 0, 0, col_shift=qcol)
        if root_pose is not None:
            assert not return_jacobian, "the Jacobian is expressed in the chain root frame: ask for it without root_pose"
            rp = (root_pose.raw_pose if isinstance(root_pose, Pose) else root_pose).to(device=q0.device, dtype=torch.float32)
            rp = rp.expand(B, 7).contiguous()
            L.check(self._lib.msik_fk_world(self._chain, C.byref(cfg), qptr, None, L.ptr(rp), L.ptr(pose), _stream()))
            return Pose(pose)
        L.check(self._lib.msik_fk(self._chain, C.byref(cfg), qptr, None, L.ptr(pose), L.ptr(jac), _stream()))
        p = Pose(pose)
        return (p, jac) if return_jacobian else p

    # ------------------------------------------------------------------ host-buffer entry
    def delta_step_host(self, h_qpos: torch.Tensor, h_action: torch.Tensor, h_out: Optional[torch.Tensor] = None, *,
                        action_offset: int = 0, pos_only: bool = False, normalize: Optional[tuple] = None,
                        chunk_envs: int = 1 << 18) -> torch.Tensor:
        """Path A (``q0 + pinv(J) a``, kinematics.py:156-172) for HOST buffers (pinned
        recommended): chunks the batch and overlaps host->device copies, the kernel and the
        device->host copy of the joint targets (libmsik ``msik_delta_step_host``).

        h_qpos (B, Nmax) fp32 host rows in PhysX layout, h_action (B, A) fp32 host flat action,
        ``normalize=(pos_low, pos_high, rot_scale)`` applies the controller's clip + scale
        (pd_ee_pose.py:197-210) in-kernel, ``None`` takes the action as already processed.
        Returns h_out (B, c) fp32 host joint targets.  Synchronous.
        """
        for t in (h_qpos, h_action):
            assert (not t.is_cuda) and t.dtype == torch.float32 and t.dim() == 2 and t.is_contiguous(), "host fp32 rows"
        B = h_qpos.shape[0]
        assert h_action.shape[0] == B
        if h_out is None:
            h_out = torch.empty((B, self.n_ctrl), dtype=torch.float32).pin_memory()
        assert h_out.shape == (B, self.n_ctrl) and h_out.is_contiguous() and h_out.dtype == torch.float32
        key = (min(chunk_envs, max(B, 1)), h_qpos.shape[1], h_action.shape[1])
        pipe = getattr(self, "_pipes", None)
        if pipe is None:
            self._pipes = pipe = {}
        if key not in pipe:
            h = C.c_void_p()
            L.check(self._lib.msik_pipeline_create(key[0], key[1], key[2], C.byref(h)))
            pipe[key] = h
        flags = L.F_USE_DELTA | L.F_NO_DRIVE | (L.F_POS_ONLY if pos_only else 0)
        low = high = (0.0, 0.0, 0.0)
        rot = 1.0
        if normalize is not None:
            low, high, rot = normalize
            flags |= L.F_NORMALIZE
        cfg = self.step_cfg(B, h_qpos.shape[1], h_qpos.shape[1], h_action.shape[1], action_offset, flags,
                            pos_low=low, pos_high=high, rot_scale=rot)
        L.check(self._lib.msik_delta_step_host(pipe[key], self._chain, C.byref(cfg), h_qpos.data_ptr(),
                                               h_action.data_ptr(), h_out.data_ptr()))
        return h_out
