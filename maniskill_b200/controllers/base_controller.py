"""Controller API contract (mirror of mani_skill/agents/controllers/base_controller.py).

Same constructor, attributes and methods as the reference (``__init__`` :42-72,
``_preprocess_action`` :120-133, ``CombinedController.set_action`` :287-300) so that
``BaseAgent.set_action`` / ``BaseEnv.step`` call into it unchanged; the arithmetic of
the EE controllers is in libmsik.so.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List

import numpy as np
import torch

from maniskill_b200 import spaces
from maniskill_b200.structs.pose import _to_tensor


def get_joints_by_names(articulation, joint_names: List[str]):
    """agents/utils.py:27-31."""
    joints = articulation.get_active_joints()
    names = [j.name for j in joints]
    return [joints[names.index(n)] for n in joint_names]


def get_active_joint_indices(articulation, joint_names: List[str]) -> torch.Tensor:
    """agents/utils.py:14-25 (int32 tensor)."""
    names = [j.name for j in articulation.get_active_joints()]
    return torch.tensor([names.index(n) for n in joint_names], dtype=torch.int32)


def flatten_action_spaces(action_spaces: Dict[str, "spaces.Box"]):
    """agents/utils.py:34-62: concatenate Box spaces, return (Box, {uid: (start, end)})."""
    lows, highs, mapping = [], [], dict()
    offset = 0
    for uid, sp in action_spaces.items():
        assert len(sp.shape) == 1, (uid, sp.shape)
        lows.append(sp.low)
        highs.append(sp.high)
        mapping[uid] = (offset, offset + sp.shape[0])
        offset += sp.shape[0]
    flat = spaces.Box(np.hstack(lows), np.hstack(highs), dtype=np.float32)
    return flat, mapping


class BaseController:
    joints: list
    active_joint_indices: torch.Tensor
    action_space: "spaces.Box"
    single_action_space: "spaces.Box"
    _original_single_action_space: "spaces.Box"

    def __init__(self, config: "ControllerConfig", articulation, control_freq: int, sim_freq: int = None, scene=None):
        self.config = config
        self.articulation = articulation
        self._control_freq = control_freq
        self.scene = scene
        if sim_freq is None:  # infer from scene
            sim_freq = round(1.0 / self.articulation.px.timestep)
        self._sim_steps = sim_freq // control_freq
        self._initialize_joints()
        self._initialize_action_space()
        self._normalize_action = getattr(self.config, "normalize_action", False)
        if self._normalize_action:
            self._clip_and_scale_action_space()
        self.action_space = self.single_action_space
        if self.scene.num_envs > 1:
            self.action_space = spaces.batch_space(self.single_action_space, n=self.scene.num_envs)

    @property
    def device(self):
        return self.articulation.device

    def _initialize_joints(self):
        joint_names = self.config.joint_names
        try:
            self.joints = get_joints_by_names(self.articulation, joint_names)
            self.active_joint_indices = get_active_joint_indices(self.articulation, joint_names).to(self.device)
        except Exception as err:
            print("Encounter error when parsing joint names.")
            print("Joint names of the articulation", [x.name for x in self.articulation.get_active_joints()])
            print("Joint names of the controller", joint_names)
            raise err

    def _initialize_action_space(self):
        raise NotImplementedError

    @property
    def control_freq(self):
        return self._control_freq

    @property
    def qpos(self):
        return self.articulation.get_qpos()[..., self.active_joint_indices.long()]

    @property
    def qvel(self):
        return self.articulation.get_qvel()[..., self.active_joint_indices.long()]

    def set_drive_property(self):
        raise NotImplementedError

    def reset(self):
        pass

    def _stepped_by_lead(self):
        """Bookkeeping of a ``set_action`` whose arithmetic another controller's launch did (CombinedController fusion)."""
        self._step = 0

    def capture(self, action: torch.Tensor):
        """CUDA-graph-backed ``set_action`` for FIXED buffers: records one ``set_action(action)`` (reading ``action``'s
        storage and the PhysX buffers in place) and returns ``replay()``.  The caller refills ``action`` in place before
        each replay; the controller's lazily materialised state is invalidated exactly as by ``set_action``.
        Not in the reference (its set_action is ~150 eager launches); here a step is one kernel node."""
        assert isinstance(action, torch.Tensor) and action.is_cuda, "capture() needs the CUDA action buffer the replays will read"
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self.set_action(action)  # builds every cached pack / plan outside the capture
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self.set_action(action)
        self._captured = (graph, action)  # keep both alive
        replay = graph.replay
        invalidate = getattr(self, "_after_replay", None)
        if invalidate is None:
            return replay

        def run():
            replay()
            invalidate()

        return run

    def _check_action_shape(self, action):
        if self.scene.num_envs > 1:
            action_dim = self.action_space.shape[1]
        else:
            action_dim = self.action_space.shape[0]
        assert tuple(action.shape) == (self.scene.num_envs, action_dim), (tuple(action.shape), action_dim)

    def _preprocess_action(self, action):
        self._check_action_shape(action)
        if self._normalize_action:
            action = self._clip_and_scale_action(action)
        return action

    def set_action(self, action):
        raise NotImplementedError

    def before_simulation_step(self):
        pass

    def get_state(self) -> dict:
        return {}

    def set_state(self, state: dict):
        pass

    def _clip_and_scale_action_space(self):
        self._original_single_action_space = self.single_action_space
        self.single_action_space = spaces.normalize_action_space(self._original_single_action_space)
        low, high = self._original_single_action_space.low, self._original_single_action_space.high
        self.action_space_low = _to_tensor(low, device=self.device)
        self.action_space_high = _to_tensor(high, device=self.device)

    def _clip_and_scale_action(self, action):
        # gym_utils.clip_and_scale_action, utils/gym_utils.py:102-105
        action = torch.clip(action, -1, 1)
        return 0.5 * (self.action_space_high + self.action_space_low) + 0.5 * (
            self.action_space_high - self.action_space_low
        ) * action


@dataclass
class ControllerConfig:
    joint_names: List[str]
    controller_cls = BaseController


class DictController(BaseController):
    def __init__(self, configs: Dict[str, ControllerConfig], articulation, control_freq: int, sim_freq: int = None,
                 scene=None):
        self.scene = scene
        self.configs = configs
        self.articulation = articulation
        self._control_freq = control_freq
        self.controllers: Dict[str, BaseController] = dict()
        for uid, config in configs.items():
            self.controllers[uid] = config.controller_cls(config, articulation, control_freq, sim_freq=sim_freq, scene=scene)
        self._initialize_action_space()
        self._initialize_joints()
        self.action_space = self.single_action_space
        if self.scene.num_envs > 1 and isinstance(self.single_action_space, spaces.Box):
            self.action_space = spaces.batch_space(self.single_action_space, n=self.scene.num_envs)

    def before_simulation_step(self):
        for controller in self.controllers.values():
            controller.before_simulation_step()

    def _initialize_action_space(self):
        self.single_action_space = {uid: c.single_action_space for uid, c in self.controllers.items()}

    def _initialize_joints(self):
        self.joints = []
        self.active_joint_indices = []
        for controller in self.controllers.values():
            self.joints.extend(controller.joints)
            self.active_joint_indices.extend(controller.active_joint_indices)

    def set_drive_property(self):
        for controller in self.controllers.values():
            controller.set_drive_property()

    def reset(self):
        for controller in self.controllers.values():
            controller.reset()

    def set_action(self, action: Dict[str, torch.Tensor]):
        for uid, controller in self.controllers.items():
            controller.set_action(action[uid])

    def get_state(self) -> dict:
        states = {}
        for uid, controller in self.controllers.items():
            state = controller.get_state()
            if len(state) > 0:
                states[uid] = state
        return states

    def set_state(self, state: dict):
        for uid, controller in self.controllers.items():
            controller.set_state(state.get(uid))


class CombinedController(DictController):
    """Flat view of several controllers: slices ``action[:, start:end]`` per controller
    (base_controller.py:278-300).  EE controllers receive the FULL flat action plus their
    column offset, so the kernel reads its slice in place (no slicing copy)."""

    def __init__(self, configs, articulation, control_freq: int, sim_freq: int = None, scene=None):
        super().__init__(configs, articulation, control_freq, sim_freq=sim_freq, scene=scene)
        if torch.device(articulation.device).type == "cuda":
            self._fusion_plan()  # built once, outside any CUDA-graph capture (it reads index tensors)

    def _initialize_action_space(self):
        super()._initialize_action_space()
        self.single_action_space, self.action_mapping = flatten_action_spaces(self.single_action_space)

    def set_action(self, action: torch.Tensor):
        # stock fused configuration: the whole dict is the lead EE controller's ONE launch (argument pack prebuilt, see
        # PDEEPosController.fast_step) plus two attribute writes per sibling
        prog = self._prog
        if prog is False:
            prog = self._program()
        if prog is not None and self.fuse_siblings and prog(action):
            return
        self._set_action_general(action)

    _prog = False  # not built yet

    def _program(self):
        """``step(action) -> handled`` when EVERY controller of the dict is served by the lead's launch, else None.
        Built once (after the fusion plan)."""
        if not self.fuse_siblings:
            return None
        if self._prog is not False:
            return self._prog
        self._prog = None
        if torch.device(self.articulation.device).type != "cuda":
            return None
        plan = self._fusion_plan()
        if plan is None or set(self.controllers) != plan["siblings"] | {plan["lead"]}:
            return None
        lead = self.controllers[plan["lead"]]
        if not hasattr(lead, "fast_step"):
            return None
        start, end = self.action_mapping[plan["lead"]]
        adim = self.action_space.shape[1] if self.scene.num_envs > 1 else self.action_space.shape[0]
        self._prog = lead.fast_step(start, end, aux=plan["aux"], vel=plan["vel"], adim=adim,
                                    siblings=[self.controllers[u] for u in plan["siblings"]], dual=plan["dual"])
        return self._prog

    def _after_replay(self):
        plan = self._fusion_plan()
        fused = plan["siblings"] if plan is not None else ()
        for uid, c in self.controllers.items():
            f = getattr(c, "_after_replay", None)
            if f is not None:
                f()
            else:
                c._step = 0
                if uid in fused:
                    c._stepped_by_lead()  # its columns were written by the lead's launch

    def _set_action_general(self, action: torch.Tensor):
        if self.scene.num_envs > 1:
            action_dim = self.action_space.shape[1]
        else:
            action_dim = self.action_space.shape[0]
        assert tuple(action.shape) == (self.scene.num_envs, action_dim), (
            f"Received action of shape {tuple(action.shape)} but expected shape ({self.scene.num_envs}, {action_dim})"
        )
        plan = self._fusion_plan() if isinstance(action, torch.Tensor) and action.is_cuda else None
        for uid, controller in self.controllers.items():
            start, end = self.action_mapping[uid]
            if plan is not None and uid in plan["siblings"]:
                if plan["dual"] is None or controller is not plan["dual"][0]:  # (the second arm's state is the lead's business)
                    controller._stepped_by_lead()  # its columns are written by the lead's launch
                continue
            fused = getattr(controller, "set_action_flat", None)
            if fused is not None and isinstance(action, torch.Tensor):
                if plan is not None and uid == plan["lead"]:
                    fused(action, start, end, aux=plan["aux"], vel=plan["vel"], dual=plan["dual"])
                else:
                    fused(action, start, end)
            else:
                controller.set_action(action[:, start:end])

    fuse_siblings = True

    def _fusion_plan(self):
        """One launch for the whole CombinedController where possible: the EE controller on the
        one-step delta path (lead) also writes the drive targets of its sibling joint-position
        controllers (no use_target / interpolate), see ``msik_aux_cfg`` in include/msik.h.
        Built once; ``None`` when there is nothing to fuse."""
        if not self.fuse_siblings:
            return None
        plan = getattr(self, "_plan", False)
        if plan is not False:
            return plan
        from maniskill_b200 import _lib as L
        from .pd_ee_pose import PDEEPosController
        from .pd_joint_pos import PDJointPosController, PDJointPosMimicController

        lead = None
        for uid, c in self.controllers.items():
            if isinstance(c, PDEEPosController) and c.config.use_delta and not c.config.use_target and not c.config.interpolate:
                lead = uid
                break
        sib = []
        if lead is not None:
            for uid, c in self.controllers.items():
                if (type(c) in (PDJointPosController, PDJointPosMimicController)
                        and not c.config.use_target and not c.config.interpolate and c.articulation is self.articulation):
                    sib.append(uid)
        # one velocity controller of the same articulation rides along (Fetch: base = pd_base_vel)
        from .pd_joint_vel import PDJointVelController

        vel_uid = None
        if lead is not None and not self.controllers[lead].config.interpolate:
            for uid, c in self.controllers.items():
                if isinstance(c, PDJointVelController) and c.articulation is self.articulation and len(c.joints) <= L.MSIK_MAX_AUX:
                    vel_uid = uid
                    break
        # a second EE controller of the same articulation on the same path (the fork's Realman robot: right and left arm,
        # mobile_realman.py:436-443) rides in the lead's launch: msik_delta_step_dual
        lead2 = None
        if lead is not None:
            lc = self.controllers[lead]
            for uid, c in self.controllers.items():
                if (uid != lead and isinstance(c, PDEEPosController) and c.config.use_delta and not c.config.use_target
                        and not c.config.interpolate and c.articulation is self.articulation and c._pos_only == lc._pos_only):
                    lead2 = uid
                    break
        if lead is None or (not sib and vel_uid is None and lead2 is None):
            self._plan = None
            return None
        aux = L.AuxCfg()
        k = 0
        for uid in sib:
            c = self.controllers[uid]
            start, end = self.action_mapping[uid]
            cols = [int(x) for x in c.active_joint_indices.tolist()]
            adim = end - start
            if c._normalize_action:
                lo, hi = c.action_space_low, c.action_space_high
                mid = (0.5 * (hi + lo)).float().cpu().numpy()  # same fp32 roundings as _clip_and_scale_action
                half = (0.5 * (hi - lo)).float().cpu().numpy()
            for j, col in enumerate(cols):
                if k >= L.MSIK_MAX_AUX:
                    self._plan = None
                    return None
                a = j if adim == len(cols) else 0  # mimic joints share one action column
                aux.col[k] = col
                aux.action_col[k] = start + a
                aux.flags[k] = (L.AUX_NORMALIZE if c._normalize_action else 0) | (L.AUX_DELTA if c.config.use_delta else 0)
                aux.mid[k] = float(mid[a]) if c._normalize_action else 0.0
                aux.half[k] = float(half[a]) if c._normalize_action else 0.0
                k += 1
        aux.n_cols = k
        vel = (self.controllers[vel_uid], self.action_mapping[vel_uid][0]) if vel_uid is not None else None
        dual = (self.controllers[lead2], self.action_mapping[lead2][0]) if lead2 is not None else None
        self._plan = {"lead": lead, "siblings": set(sib) | ({vel_uid} if vel_uid is not None else set()) | ({lead2} if lead2 else set()),
                      "aux": aux if k else None, "vel": vel, "dual": dual}
        return self._plan

    def to_action_dict(self, action):
        action_dim = self.action_space.shape[0] if self.scene.num_envs == 1 else self.action_space.shape[1]
        assert action.shape == (action_dim,), (action.shape, action_dim)
        return {uid: action[s:e] for uid, (s, e) in self.action_mapping.items()}

    def from_action_dict(self, action_dict: dict):
        return torch.hstack([action_dict[uid] for uid in self.controllers])
