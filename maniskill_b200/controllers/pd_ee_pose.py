"""PD end-effector position / pose controllers on the fused sm_100a path.

Drop-in for mani_skill/agents/controllers/pd_ee_pose.py: same classes, config
dataclasses, attributes and methods (``set_action`` :91-115, ``reset`` :66-74,
``compute_target_pose`` :76-89 / :212-238, ``get_state`` / ``set_state`` :117-127,
``ee_pose_at_base`` :61-64).  ``set_action`` is ONE kernel launch that reads the
PhysX qpos buffer and the flat action in place and writes the PhysX drive-target
buffer in place (libmsik.so: msik_delta_step / msik_target_ik).

State bookkeeping (``state_mode``):
  "lazy"  (default) the kernel writes only the drive targets; ``_target_qpos`` is
          gathered from the drive-target buffer when somebody reads it, ``_start_qpos``
          is read from the qpos buffer AT READ TIME (the reference snapshots it in
          ``set_action``; identical until the simulator steps), and on the delta path
          (``use_delta and not use_target``) ``_target_pose`` reads ``None`` after a step
          because nothing on that path consumes it.
  "eager" every ``set_action`` also materialises ``_start_qpos``, ``_target_qpos`` and
          ``_target_pose`` exactly like the reference does (2 extra small launches on
          the delta path).
``interpolate`` and ``use_target`` always keep the state they need.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence, Union

import numpy as np
import torch

from maniskill_b200 import _lib as L
from maniskill_b200 import spaces
from maniskill_b200.kinematics import Kinematics, _stream
from maniskill_b200.structs.pose import Pose

from .base_controller import ControllerConfig
from .pd_joint_pos import PDJointPosController


class PDEEPosController(PDJointPosController):
    """The PD EE Position controller (GPU sim; one robot per sub-scene)."""

    config: "PDEEPosControllerConfig"
    _pos_only = True
    state_mode = "lazy"

    # ------------------------------------------------------------------ init
    def _check_gpu_sim_works(self):
        assert (
            self.config.frame == "root_translation"
        ), "currently only translation in the root frame for EE control is supported in GPU sim"

    def _initialize_joints(self):
        self.initial_qpos = None
        super()._initialize_joints()
        if torch.device(self.device).type == "cuda":
            self._check_gpu_sim_works()
        self.kinematics = Kinematics(self.config.urdf_path, self.config.ee_link, self.articulation, self.active_joint_indices)
        self.ee_link = self.kinematics.end_link
        self._lib = L.load()
        self._tp_buf: Optional[torch.Tensor] = None  # (B,7) target pose state
        self._tq_buf: Optional[torch.Tensor] = None  # (B,c) dense target qpos
        self._sq_buf: Optional[torch.Tensor] = None  # (B,c) dense start qpos
        self._tq_valid = self._sq_valid = self._tp_valid = False
        self._bufs = None
        self._cfg_cache = dict()
        self._step_handles = []
        self._step = 0

    def _initialize_action_space(self):
        low = np.float32(np.broadcast_to(self.config.pos_lower, 3))
        high = np.float32(np.broadcast_to(self.config.pos_upper, 3))
        self.single_action_space = spaces.Box(low, high, dtype=np.float32)

    # ------------------------------------------------------------------ PhysX buffers
    def _physx(self):
        """(qpos, target_qpos, row_index|None) views of the PhysX GPU buffers, cached."""
        if self._bufs is None:
            px = self.articulation.px
            qpos = px.cuda_articulation_qpos.torch()
            tgt = px.cuda_articulation_target_qpos.torch()
            idx = self.articulation._data_index
            B = self.scene.num_envs
            assert idx.shape[0] == B, (idx.shape, B)
            ident = bool(torch.equal(idx.long().cpu(), torch.arange(B)))
            self._bufs = (qpos, tgt, None if ident else idx.to(torch.int32).contiguous())
        return self._bufs

    def invalidate_buffers(self):
        """Call if SAPIEN re-allocates its CUDA buffers (scene reconfiguration)."""
        self._bufs = None
        self._cfg_cache = dict()
        getattr(self, "_packs", dict()).clear()  # the fast_step closures hold this dict
        self._drop_step_handles()

    def _drop_step_handles(self):
        for h in getattr(self, "_step_handles", []):
            self._lib.msik_step_destroy(h)
        self._step_handles = []

    def __del__(self):
        try:
            self._drop_step_handles()
        except Exception:
            pass

    def _flags(self) -> int:
        c = self.config
        f = 0
        if self._normalize_action:
            f |= L.F_NORMALIZE
        if self._pos_only:
            f |= L.F_POS_ONLY
        if c.use_delta:
            f |= L.F_USE_DELTA
        if c.use_target:
            f |= L.F_USE_TARGET
        if "body_translation" in c.frame:
            f |= L.F_BODY_TRANS
        if "body_aligned_body_rotation" in c.frame:
            f |= L.F_BODY_ROT
        if c.interpolate:
            f |= L.F_NO_DRIVE
        return f

    def _step_cfg(self, action_stride: int, action_offset: int, flags: int) -> L.StepCfg:
        key = (action_stride, action_offset, flags)
        cfg = self._cfg_cache.get(key)
        if cfg is None:
            qpos, tgt, _ = self._physx()
            if self._normalize_action:
                low = self._original_single_action_space.low
                high = self._original_single_action_space.high
            else:
                low = high = np.zeros(6, np.float32)
            rot = self.config.rot_lower if getattr(self.config, "rot_lower", None) is not None else 1.0
            assert np.ndim(rot) == 0 or len(set(np.ravel(rot).tolist())) == 1, "per-axis rot_lower is not supported"
            cfg = self.kinematics.step_cfg(
                self.scene.num_envs, qpos.stride(0), tgt.stride(0), action_stride, action_offset, flags,
                pos_low=low[:3], pos_high=high[:3], rot_scale=float(np.ravel(rot)[0]),
            )
            self._cfg_cache[key] = cfg
        return cfg

    def _dense(self, name: str, cols: int) -> torch.Tensor:
        buf = getattr(self, name)
        if buf is None:
            buf = torch.empty((self.scene.num_envs, cols), dtype=torch.float32, device=self.device)
            setattr(self, name, buf)
        return buf

    # ------------------------------------------------------------------ reference attributes
    @property
    def _target_qpos(self):
        if self._tq_buf is None and not self._tq_valid and self._bufs is None:
            return None
        if not self._tq_valid:
            _, tgt, idx = self._physx()
            rows = tgt if idx is None else tgt[idx.long()]
            self._tq_buf = rows[: self.scene.num_envs, self.active_joint_indices.long()].contiguous()
            self._tq_valid = True
        return self._tq_buf

    @_target_qpos.setter
    def _target_qpos(self, v):
        self._tq_buf = v
        self._tq_valid = v is not None

    @property
    def _start_qpos(self):
        if not self._sq_valid:
            if self._bufs is None and self._sq_buf is None:
                return None
            self._sq_buf = self.qpos
            self._sq_valid = True
        return self._sq_buf

    @_start_qpos.setter
    def _start_qpos(self, v):
        self._sq_buf = v
        self._sq_valid = v is not None

    @property
    def _target_pose(self):
        if not self._tp_valid:
            return None
        return Pose(self._tp_buf)

    @_target_pose.setter
    def _target_pose(self, v):
        if v is None:
            self._tp_buf, self._tp_valid = None, False
        else:
            self._tp_buf = Pose.create(v, device=self.device).raw_pose.contiguous().float()
            self._tp_valid = True

    # ------------------------------------------------------------------ poses
    @property
    def ee_pos(self):
        return self.ee_link.pose.p

    @property
    def ee_pose(self):
        return self.ee_link.pose

    def _body_rows(self):
        """int32 rows of the root / EE link in ``cuda_rigid_body_data``, converted ONCE and kept alive on the
        controller: the reference's ``_body_data_index`` is int64 (utils/structs/base.py:104-110), so a per-call
        ``.to(torch.int32)`` would hand the launch the pointer of a temporary that the caching allocator may
        already have given to the next conversion."""
        rows = getattr(self, "_body_rows_cache", None)
        if rows is None:
            root = getattr(self.articulation, "root", None)
            root_idx = getattr(root, "_body_data_index", None) if root is not None else None
            ee_idx = getattr(self.ee_link, "_body_data_index", None)
            if root_idx is None or ee_idx is None:
                rows = (None, None)
            else:
                dev = self.device
                rows = (torch.as_tensor(root_idx, device=dev).to(torch.int32).contiguous().clone(),
                        torch.as_tensor(ee_idx, device=dev).to(torch.int32).contiguous().clone())
            self._body_rows_cache = rows
        return rows

    @property
    def ee_pose_at_base(self) -> Pose:
        """articulation.pose.inv() * ee_link.pose (pd_ee_pose.py:61-64), one launch."""
        root_idx, ee_idx = self._body_rows()
        if root_idx is None:
            return self.articulation.pose.inv() * self.ee_pose
        body = self.articulation.px.cuda_rigid_body_data.torch()
        B = self.scene.num_envs
        out = torch.empty((B, 7), dtype=torch.float32, device=self.device)
        L.check(self._lib.msik_ee_pose_at_base(L.ptr(body), body.stride(0), L.ptr(root_idx), L.ptr(ee_idx), B, L.ptr(out),
                                               _stream()))
        return Pose(out)

    # ------------------------------------------------------------------ reset
    def reset(self):
        """pd_joint_pos.py:51-66 + pd_ee_pose.py:66-74, masked by scene._reset_mask.

        Only a controller that was NEVER initialised takes every env; state that lazy mode left un-materialised
        is first re-gathered from the PhysX buffers (``_target_qpos`` / ``_start_qpos`` properties) and then
        updated under the real mask, so a partial reset never touches the envs that keep running."""
        self._step = 0
        qpos, _, idx = self._physx()
        B, c = self.scene.num_envs, self.kinematics.n_ctrl
        mask = self.scene._reset_mask
        never = self._sq_buf is None or self._tq_buf is None
        if never:
            # nothing to keep: initialise every env (pd_joint_pos.py:54-55,61-62)
            self._sq_buf = torch.empty((B, c), dtype=torch.float32, device=self.device)
            self._tq_buf = torch.empty((B, c), dtype=torch.float32, device=self.device)
            m8 = torch.ones(B, dtype=torch.uint8, device=self.device)
        else:
            if not self._tq_valid:
                _ = self._target_qpos  # gather the running envs' targets from the drive buffer
            if not self._sq_valid:
                _ = self._start_qpos
            m8 = mask.to(torch.uint8).contiguous()
        ee = self.ee_pose_at_base.raw_pose.contiguous()
        tp_ptr = ee_ptr = None
        if self._tp_buf is None or not self._tp_valid:
            # pd_ee_pose.py:68-69 (`_target_pose is None`); also a pose that lazy path-A stepping left stale
            self._tp_buf = ee.clone()
            self._tp_valid = True
        elif never:
            # target pose restored by set_state before the first reset: masked update (pd_ee_pose.py:70-74)
            self._tp_buf[mask.bool()] = ee[mask.bool()]
        else:
            tp_ptr, ee_ptr = L.ptr(self._tp_buf), L.ptr(ee)
        cfg = self._step_cfg(0, 0, 0)
        L.check(self._lib.msik_reset_masked(C.byref(cfg), L.ptr(m8), L.ptr(qpos), L.ptr(idx), ee_ptr, L.ptr(self._sq_buf),
                                            L.ptr(self._tq_buf), tp_ptr, _stream()))
        self._sq_valid = self._tq_valid = True

    # ------------------------------------------------------------------ target pose (API)
    def compute_target_pose(self, prev_ee_pose_at_base: Pose, action: torch.Tensor) -> Pose:
        """pd_ee_pose.py:76-89 / :212-238 on an already pre-processed action."""
        action = Kinematics._rows(action)
        B = action.shape[0]
        flags = self._flags() & ~(L.F_NORMALIZE | L.F_NO_DRIVE)
        action, aptr, astride, acol = Kinematics._base_view(action)
        cfg = self._step_cfg(astride, acol, flags)
        prev = None
        if self.config.use_delta:
            prev = Pose.create(prev_ee_pose_at_base, device=self.device).raw_pose.contiguous().float()
            if prev.shape[0] == 1 and B > 1:
                prev = prev.repeat(B, 1)
        out = torch.empty((B, 7), dtype=torch.float32, device=self.device)
        L.check(self._lib.msik_target_pose(C.byref(cfg), aptr, L.ptr(prev), L.ptr(out), _stream()))
        return Pose(out)

    # ------------------------------------------------------------------ set_action
    def set_action(self, action):
        own = self.__dict__.get("_own_step", False)
        if own is False:  # built on first use: the prepared step of a controller driven on its own
            nd = 3 if self._pos_only else 6
            own = self._own_step = self.fast_step(0, nd) if torch.device(self.device).type == "cuda" else None
        if own is not None and own(action):
            return
        self._check_action_shape(action)
        self.set_action_flat(action, 0, action.shape[1])

    def _after_replay(self):
        """State bookkeeping of one ``set_action`` replayed from a CUDA graph (BaseController.capture)."""
        self._step = 0
        if self.state_mode == "lazy" and not self.config.interpolate:
            self._tq_valid = self._sq_valid = False
            if self.config.use_delta and not self.config.use_target:
                self._tp_valid = False

    def _stepped_by_lead(self):
        """This controller's columns were just written by ANOTHER EE controller's launch (second arm of a dual step):
        the same bookkeeping as after its own lazy delta step."""
        self._step = 0
        self._tq_valid = self._sq_valid = self._tp_valid = False

    def _dual_cfg(self, second, stride: int) -> "L.DualCfg":
        """``msik_dual_cfg`` for ``second = (controller, action start)`` on a flat action of ``stride`` columns."""
        c2, start2 = second
        d = L.DualCfg()
        d.chain = c2.kinematics._chain
        d.cfg = c2._step_cfg(stride, start2, c2._flags())
        return d

    def fast_step(self, start: int, end: int, aux=None, vel=None, adim=None, siblings=(), dual=None):
        """``step(flat_action)`` for the stock stepping configuration (delta path, lazy state, drive targets written):
        the argument pack of ``msik_delta_step_fused_vel`` is built ONCE per action row stride -- step cfg, sibling /
        velocity cfgs, PhysX buffer pointers as ready-made ctypes objects -- so a step is the tensor checks, one dict
        lookup and the foreign call.  Anything else (numpy / strided / column-sliced actions, eager state,
        interpolate, use_target) takes ``set_action_flat``.  At RL-typical env counts (4096) the host issue path IS
        the step time."""
        conf = self.config
        if not (conf.use_delta and not conf.use_target and not conf.interpolate):
            return None
        adim = (end - start) if adim is None else adim  # columns of the flat action (CombinedController: the whole dict)
        siblings = tuple(siblings)  # controllers whose columns this launch writes: marked stale with the step
        launch = self._lib.msik_step_launch
        packs = self._packs = getattr(self, "_packs", dict())
        f32 = torch.float32
        B = self.scene.num_envs
        raw_stream = torch._C._cuda_getCurrentRawStream
        dev_index = torch.device(self.device).index
        if dev_index is None:  # "cuda" without an index = the process's current device (one process per GPU)
            dev_index = torch.cuda.current_device()
        check = L.check
        ctrl = self
        import weakref

        last = [lambda: None, None]  # (weak reference to the tensor whose layout was checked last, its handle)

        def build(stride: int):
            qpos, tgt, idx = ctrl._physx()
            cfg = ctrl._step_cfg(stride, start, ctrl._flags())
            vcfg = qvel = None
            if vel is not None:
                vctrl, vstart = vel
                vcfg = vctrl._vel_cfg(stride, vstart)
                qvel = ctrl.articulation.px.cuda_articulation_target_qvel.torch()
            h = C.c_void_p()
            dcfg = ctrl._dual_cfg(dual, stride) if dual is not None else None
            check(ctrl._lib.msik_step_create_dual(ctrl.kinematics._chain, C.byref(cfg), C.byref(dcfg) if dcfg is not None else None,
                                                  C.byref(aux) if aux is not None else None,
                                                  C.byref(vcfg) if vcfg is not None else None, L.ptr(qpos), L.ptr(idx), L.ptr(tgt),
                                                  L.ptr(qvel), C.byref(h)))
            ctrl._step_handles.append(h)
            return h, (qpos, tgt, idx, qvel)  # the handle refers to these buffers: keep them alive with it

        key = (adim, start, id(aux), id(vel), id(dual))

        def step(a):
            """True when the step was issued here; False -> the caller takes the general path."""
            if ctrl.state_mode != "lazy" or (dual is not None and dual[0].state_mode != "lazy"):
                return False
            if a is last[0]() and a.numel() == B * adim:
                h = last[1]  # same tensor object as the step before, not resized since: its layout was checked then
            else:
                if type(a) is not torch.Tensor or a.dtype is not f32 or not a.is_cuda or not a.is_contiguous():
                    return False
                sh = a.shape
                if len(sh) != 2 or sh[0] != B or sh[1] != adim:
                    return False
                pk = packs.get(key)
                if pk is None:
                    pk = packs[key] = build(adim)
                h = pk[0]
                last[0], last[1] = weakref.ref(a), h
            rc = launch(h, a.data_ptr(), raw_stream(dev_index))
            if rc:
                check(rc)
            ctrl._step = 0
            ctrl._tq_valid = ctrl._sq_valid = ctrl._tp_valid = False
            for c in siblings:
                c._stepped_by_lead()
            return True

        return step

    def set_action_flat(self, flat_action: torch.Tensor, start: int, end: int, aux=None, vel=None, dual=None):
        """Fused ``set_action`` on columns [start, end) of the flat action tensor
        (CombinedController passes the whole tensor: no slice copy).  ``aux`` (``_lib.AuxCfg``):
        sibling joint-position controllers whose drive targets the same launch writes; ``vel`` = (controller,
        action start): the velocity controller of the same CombinedController, written by the same launch too."""
        nd = 3 if self._pos_only else 6
        assert end - start == nd, (start, end, nd)
        a = flat_action
        if not (isinstance(a, torch.Tensor) and a.is_cuda and a.dtype == torch.float32 and a.dim() == 2 and a.stride(1) == 1):
            a = torch.as_tensor(np.asarray(a) if not isinstance(a, torch.Tensor) else a, device=self.device).float().contiguous()
        assert a.shape[0] == self.scene.num_envs, (a.shape, self.scene.num_envs)
        self._step = 0
        qpos, tgt, idx = self._physx()
        c = self.kinematics.n_ctrl
        cfgf = self._flags()
        a, aptr, astride, acol = Kinematics._base_view(a)
        cfg = self._step_cfg(astride, start + acol, cfgf)
        if aux is not None and acol:
            # the flat action is itself a column slice of a wider buffer: the kernel is handed the
            # parent row, so the sibling action columns shift with it
            shifted = L.AuxCfg.from_buffer_copy(aux)
            for k in range(aux.n_cols):
                shifted.action_col[k] = aux.action_col[k] + acol
            aux = shifted
        conf = self.config
        eager = self.state_mode == "eager"
        want_dense = eager or conf.interpolate
        tq = self._dense("_tq_buf", c) if want_dense else None
        sq = self._dense("_sq_buf", c) if want_dense else None
        drive = None if conf.interpolate else tgt

        if conf.use_delta and not conf.use_target:
            if eager:
                prev = self.ee_pose_at_base.raw_pose
                tp = self._dense("_tp_buf", 7)
                L.check(self._lib.msik_target_pose(C.byref(cfg), aptr, L.ptr(prev), L.ptr(tp), _stream()))
                self._tp_valid = True
            else:
                # lazy: path A never needs the target pose (kinematics.py:156-172 takes the action only); the pose of
                # an earlier step / reset must not be mistaken for this step's -> `_target_pose` reads None
                self._tp_valid = False
            vcfg = qvel = None
            if vel is not None:
                vctrl, vstart = vel
                vcfg = vctrl._vel_cfg(astride, vstart + acol)
                qvel = self.articulation.px.cuda_articulation_target_qvel.torch()
            fused_second = (dual is not None and not want_dense and not eager and dual[0].state_mode == "lazy"
                            and not dual[0].config.interpolate)
            if fused_second:  # both arms from one pass over the rows (msik_delta_step_dual)
                dcfg = self._dual_cfg((dual[0], dual[1] + acol), astride)
                L.check(self._lib.msik_delta_step_dual(self.kinematics._chain, C.byref(cfg), C.byref(dcfg),
                                                       C.byref(aux) if aux is not None else None,
                                                       C.byref(vcfg) if vcfg is not None else None, L.ptr(qpos), L.ptr(idx),
                                                       aptr, L.ptr(drive), L.ptr(qvel), _stream()))
                dual[0]._stepped_by_lead()
            else:
                L.check(self._lib.msik_delta_step_fused_vel(self.kinematics._chain, C.byref(cfg),
                                                            C.byref(aux) if aux is not None else None,
                                                            C.byref(vcfg) if vcfg is not None else None, L.ptr(qpos), L.ptr(idx),
                                                            aptr, L.ptr(drive), L.ptr(qvel), L.ptr(tq), L.ptr(sq), _stream()))
                if dual is not None:  # state bookkeeping the fused launch cannot serve: the second arm steps on its own
                    nd2 = 3 if dual[0]._pos_only else 6
                    dual[0].set_action_flat(flat_action, dual[1], dual[1] + nd2)
        else:
            prev = None
            if conf.use_delta:  # use_target: previous target pose is the state
                if not self._tp_valid:
                    self._tp_buf = self.ee_pose_at_base.raw_pose.contiguous()
                    self._tp_valid = True
                prev = self._tp_buf
            tp = self._dense("_tp_buf", 7)
            it = self.kinematics.iter_cfg()
            L.check(self._lib.msik_target_ik(self.kinematics._chain, C.byref(cfg), C.byref(it), L.ptr(qpos), L.ptr(idx),
                                             aptr, L.ptr(prev), L.ptr(tp), L.ptr(drive), L.ptr(tq), L.ptr(sq),
                                             None, None, _stream()))
            self._tp_valid = True
        self._tq_valid = self._sq_valid = bool(want_dense)
        if conf.interpolate:
            self._step_size = (self._tq_buf - self._sq_buf) / self._sim_steps

    def before_simulation_step(self):
        self._step += 1
        if self.config.interpolate:
            _, tgt, idx = self._physx()
            cfg = self._step_cfg(0, 0, 0)
            L.check(self._lib.msik_interp_targets(C.byref(cfg), L.ptr(self._sq_buf), L.ptr(self._step_size),
                                                  float(self._step), L.ptr(idx), L.ptr(tgt), _stream()))

    def get_state(self) -> dict:
        if self.config.use_target:
            return {"target_pose": self._target_pose.raw_pose}
        return {}

    def set_state(self, state: dict):
        if self.config.use_target:
            target_pose = state["target_pose"]
            self._target_pose = Pose.create_from_pq(target_pose[:, :3], target_pose[:, 3:])


@dataclass
class PDEEPosControllerConfig(ControllerConfig):
    pos_lower: Union[float, Sequence[float]]
    pos_upper: Union[float, Sequence[float]]
    stiffness: Union[float, Sequence[float]]
    damping: Union[float, Sequence[float]]
    force_limit: Union[float, Sequence[float]] = 1e10
    friction: Union[float, Sequence[float]] = 0.0
    ee_link: str = None
    urdf_path: str = None
    frame: str = "root_translation"
    use_delta: bool = True
    use_target: bool = False
    interpolate: bool = False
    normalize_action: bool = True
    drive_mode: Union[Sequence[str], str] = "force"
    controller_cls = PDEEPosController


class PDEEPoseController(PDEEPosController):
    config: "PDEEPoseControllerConfig"
    _pos_only = False

    def _check_gpu_sim_works(self):
        assert (
            self.config.frame == "root_translation:root_aligned_body_rotation"
        ), "currently only translation in the root frame for EE control is supported in GPU sim"

    def _initialize_action_space(self):
        low = np.float32(np.hstack([np.broadcast_to(self.config.pos_lower, 3), np.broadcast_to(self.config.rot_lower, 3)]))
        high = np.float32(np.hstack([np.broadcast_to(self.config.pos_upper, 3), np.broadcast_to(self.config.rot_upper, 3)]))
        self.single_action_space = spaces.Box(low, high, dtype=np.float32)

    def _clip_and_scale_action(self, action):
        """pd_ee_pose.py:197-210 in torch (API parity; the fused path does this in-kernel)."""
        pos = torch.clip(action[:, :3], -1, 1)
        lo, hi = self.action_space_low[:3], self.action_space_high[:3]
        pos = 0.5 * (hi + lo) + 0.5 * (hi - lo) * pos
        rot = action[:, 3:].clone()
        nrm = torch.linalg.norm(rot, dim=1)
        rot = torch.where((nrm > 1)[:, None], rot * (1 / nrm[:, None]), rot)
        return torch.hstack([pos, rot * self.config.rot_lower])


@dataclass
class PDEEPoseControllerConfig(PDEEPosControllerConfig):
    rot_lower: Union[float, Sequence[float]] = None
    rot_upper: Union[float, Sequence[float]] = None
    stiffness: Union[float, Sequence[float]] = None
    damping: Union[float, Sequence[float]] = None
    force_limit: Union[float, Sequence[float]] = 1e10
    friction: Union[float, Sequence[float]] = 0.0
    frame: str = "root_translation:root_aligned_body_rotation"
    controller_cls = PDEEPoseController
