#!/usr/bin/env python3
"""bench.py -- IK controller steps/s (envs*steps/s) of the batched EE-controller hot path.

One "step" = one ``PDEEPoseController.set_action`` (gather qpos -> clip/scale action -> FK ->
Jacobian -> min-norm solve -> scatter drive targets) over this rank's env shard: ONE launch
of the fused sm_100a kernel (libmsik.so).  Workload: BASELINE.json configs[4] point
"synthetic Panda pd_ee_delta_pose, 2^20 envs per GPU" (the >=1M-env size north_star quotes the
HBM-roofline target on); ``--envs 4096`` gives configs[1].  Env-sharded weak scaling: every
rank owns ``--envs`` envs, there is no collective inside the step (NCCL only all-gathers the
per-shard statistics after the timed region).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs B] [--impl ours|reference]

Prints ONE JSON line on rank 0 (see DESIGN.md "Measurement" for every key).
``--impl reference`` times the reference's CPU arithmetic for the path on the host cores (oracle/ik_oracle.c, fp64:
neither pytorch_kinematics nor sapien is installable): headline = the path-A port, the algorithm the GPU arm runs;
CLIK, what sim_backend=cpu really runs, and the single-thread one-env-per-call figures ride along.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

URDF = os.path.join(ROOT, "maniskill_b200", "assets", "robots", "panda_v2_kin.urdf")
FETCH_URDF = os.path.join(ROOT, "maniskill_b200", "assets", "robots", "fetch_kin.urdf")
EE = "panda_hand_tcp"
ARM = [f"panda_joint{i}" for i in range(1, 8)]
FETCH_ARM = ["shoulder_pan_joint", "shoulder_lift_joint", "upperarm_roll_joint", "elbow_flex_joint", "forearm_roll_joint",
             "wrist_flex_joint", "wrist_roll_joint"]
PANDA_REST = np.array([0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4, 0.04, 0.04])  # panda.py:35-48
FETCH_REST = np.array([0, 0, 0, 0.386, 0, 0, 0, -np.pi / 4, 0, np.pi / 4, 0, np.pi / 3, 0, 0.015, 0.015])  # fetch.py:44-48
BYTES_PER_SOLVE = {"panda_pose": 80, "panda_pos": 68, "fetch_pose": 96}  # qpos in + action in + targets out
IK_FLOP_PER_ENV_ITER = 2 * 323 + 119 + 29  # loop body of target_ik_kernel<panda_tcp>: 323 FFMA, 119 FMUL, 29 FADD (profiles/r02a_target_ik_stall_by_opcode.txt)
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4: one FFMA per lane per cycle at the max SM clock
L2_BYTES = 126 * 1024 * 1024
METRIC = "IK controller steps/sec (envs*steps/s)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while a timed region is active."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, index: int):
        self.samples, self.reasons, self.active, self._stop = [], set(), False, False
        self.max_mhz = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        while not self._stop:
            if self.active and self.nv is not None:
                try:
                    self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                    r = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                    for k, bit in {**self.BAD, **self.NOTE}.items():
                        if r & bit:
                            self.reasons.add(k)
                except Exception:
                    pass
            time.sleep(0.002)

    def summary(self):
        self._stop = True
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ----------------------------------------------------------------------------- workload
class Group:
    """One env group: shim articulation with PhysX-layout buffers + the controller on it."""

    def __init__(self, robot: str, B: int, seed: int, mode: str = "pd_ee_delta_pose", fused: bool = False):
        import torch
        from maniskill_b200.controllers import PDEEPosControllerConfig, PDEEPoseControllerConfig
        from maniskill_b200.shim import FakeArticulation

        urdf, ee, arm, rest = (URDF, EE, ARM, PANDA_REST) if robot == "panda" else (FETCH_URDF, "gripper_link", FETCH_ARM, FETCH_REST)
        self.art = FakeArticulation(urdf, B, ee_link_names=[ee])
        kw = dict(joint_names=arm, pos_lower=-0.1, pos_upper=0.1, stiffness=1e3, damping=1e2, force_limit=100, ee_link=ee,
                  urdf_path=urdf)
        if mode == "pd_ee_delta_pos":
            cfg = PDEEPosControllerConfig(**kw)
        else:
            cfg = PDEEPoseControllerConfig(rot_lower=-0.1, rot_upper=0.1, **kw)
            if mode == "pd_ee_pose":  # absolute target, iterative IK (panda.py:118-121)
                cfg.use_delta = False
                cfg.normalize_action = False
        self.ctrl = cfg.controller_cls(cfg, self.art, control_freq=20, sim_freq=100, scene=self.art.scene)
        self.combined = None
        if fused and robot == "panda" and mode == "pd_ee_delta_pose":
            # the stock Panda controller dict (panda.py:103-141,192-194): arm + mimic gripper in one launch
            from maniskill_b200.controllers import CombinedController, PDJointPosMimicControllerConfig

            grip = PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01,
                                                   upper=0.04, stiffness=1e3, damping=1e2, force_limit=100)
            self.combined = CombinedController(dict(arm=cfg, gripper=grip), self.art, control_freq=20, sim_freq=100,
                                               scene=self.art.scene)
            self.ctrl = self.combined.controllers["arm"]
        g = torch.Generator(device="cuda").manual_seed(2022 + seed)  # sapien_env.py:305 default seed
        rest_t = torch.tensor(rest, dtype=torch.float32, device="cuda")
        qpos = self.art.px.cuda_articulation_qpos.torch()
        # keyframe `rest` + N(0, 0.02) (robot_init_qpos_noise, pick_cube.py:44)
        qpos.copy_(rest_t[None] + 0.02 * torch.randn(qpos.shape, generator=g, device="cuda"))
        adim = self.ctrl.single_action_space.shape[0]
        self.adim = adim
        # flat action of the CombinedController: arm slice + 1 gripper column (panda.py:192-194)
        self.action = 2 * torch.rand((B, adim + 1), generator=g, device="cuda") - 1  # gpu_sim.py:97-100
        if mode == "pd_ee_pose":
            # reachable absolute targets (PegInsertionSide-style pd_ee_pose): FK of a perturbed state,
            # written as (xyz, euler XYZ) -- R = Rx(a) Ry(b) Rz(c), rotation_conversions.py:197-223
            q2 = qpos.clone()
            q2[:, :7] += 0.3 * (torch.rand((B, 7), generator=g, device="cuda") - 0.5)
            T = self.ctrl.kinematics.forward_kinematics(q2).to_transformation_matrix()
            R = T[:, :3, :3]
            eul = torch.stack([torch.atan2(-R[:, 1, 2], R[:, 2, 2]), torch.asin(R[:, 0, 2].clamp(-1, 1)),
                               torch.atan2(-R[:, 0, 1], R[:, 0, 0])], 1)
            self.action[:, :3] = T[:, :3, 3]
            self.action[:, 3:6] = eul
        (self.combined or self.ctrl).reset()

    def step(self):
        if self.combined is not None:
            self.combined.set_action(self.action)
        else:
            self.ctrl.set_action_flat(self.action, 0, self.adim)

    def set_bytes(self):
        return sum(t.numel() * 4 for t in (self.art.px.cuda_articulation_qpos.torch(),
                                           self.art.px.cuda_articulation_target_qpos.torch(), self.action))


class PresetGroup:
    """One env group of a stock controller dict from maniskill_b200.presets on the PhysX-layout shim."""

    def __init__(self, cfgs, urdf: str, ee_links, B: int, seed: int):
        import torch
        from maniskill_b200.controllers import CombinedController
        from maniskill_b200.shim import FakeArticulation
        from maniskill_b200.structs.pose import Pose

        self.art = FakeArticulation(urdf, B, ee_link_names=list(ee_links))
        self.combined = CombinedController(cfgs, self.art, control_freq=20, sim_freq=100, scene=self.art.scene)
        g = torch.Generator(device="cuda").manual_seed(2022 + seed)
        qpos = self.art.px.cuda_articulation_qpos.torch()
        qpos.copy_(1.5 * (2 * torch.rand(qpos.shape, generator=g, device="cuda") - 1))
        root = Pose.create(torch.tensor([[0.0, 0, 0, 1, 0, 0, 0]], device="cuda").repeat(B, 1))
        self.art.set_root_pose(root)
        for c in self.combined.controllers.values():
            if hasattr(c, "kinematics"):
                self.art.set_link_pose(c.ee_link.name, root * c.kinematics.forward_kinematics(self.art.get_qpos()))
        self.combined.reset()
        self.action = 2 * torch.rand((B, self.combined.single_action_space.shape[0]), generator=g, device="cuda") - 1

    def step(self):
        self.combined.set_action(self.action)

    def set_bytes(self):
        return sum(t.numel() * 4 for t in (self.art.px.cuda_articulation_qpos.torch(),
                                           self.art.px.cuda_articulation_target_qpos.torch(), self.action))


FUSED = True  # CombinedController(arm + gripper), sibling columns fused into the launch (--arm-only: off)


SEED_BASE = 0  # rank * 64: every rank's shard holds its own states / actions


def make_groups(robot: str, B: int, mode: str = "pd_ee_delta_pose"):
    """Ring of env groups whose buffers together exceed 3x L2 (so no step finds its inputs in
    L2), or a single group + explicit L2 flush between steps when a group is small."""
    g0 = Group(robot, B, SEED_BASE, mode, FUSED)
    per = g0.set_bytes()
    ring = -(-3 * L2_BYTES // per)
    if ring <= 8:
        return [g0] + [Group(robot, B, SEED_BASE + i, mode, FUSED) for i in range(1, max(ring, 2))], "ring"
    return [g0], "flush"


def time_steps(groups, l2mode: str, steps: int, warmup: int, sampler=None):
    """Device time of exactly `steps` controller steps (CUDA events on the launching stream).
    ring : steps issued back to back, rotating over the groups (inputs > L2), one event pair.
    flush: an L2-sized buffer is rewritten between steps, outside each step's event pair."""
    import torch

    for i in range(max(warmup, 3)):
        groups[i % len(groups)].step()
    torch.cuda.synchronize()
    if sampler:
        sampler.active = True
    if l2mode == "ring":
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            groups[i % len(groups)].step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        flush = torch.empty(2 * L2_BYTES // 4, dtype=torch.float32, device="cuda")
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for i in range(steps):
            flush.zero_()
            ev[i][0].record()
            groups[0].step()
            ev[i][1].record()
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in ev)
    if sampler:
        sampler.active = False
    return ms


def time_graph(groups, steps: int, warmup: int, sampler=None):
    """Same steps captured once in a CUDA graph (one node per step, rotating over the ring) and
    replayed: removes the Python/ctypes issue overhead so the device time is the kernels'."""
    import torch

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for i in range(max(warmup, 3, len(groups))):
            groups[i % len(groups)].step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for i in range(steps):  # exactly `steps` kernel nodes
            groups[i % len(groups)].step()
    graph.replay()  # warm replay (uploads the graph)
    torch.cuda.synchronize()
    if sampler:
        sampler.active = True
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize()
    if sampler:
        sampler.active = False
    return e0.elapsed_time(e1)


# ----------------------------------------------------------------------------- CPU arms
def cpu_inputs(B: int, seed: int = 2022):
    """Same distribution as the GPU workload, fp64, plus the target poses the reference's CPU
    path hands to CLIK: ee_pose_at_base (= FK(q)) composed with the processed action
    (pd_ee_pose.py:96-109)."""
    from oracle import oracle as O

    fr = O.load_chain(URDF, EE)
    rng = np.random.default_rng(seed)
    q = PANDA_REST[None, :7] + rng.normal(0, 0.02, (B, 7))
    a = O.pose_clip_and_scale(rng.uniform(-1, 1, (B, 6)), -0.1, 0.1, -0.1)
    _, R, p = O.jacobian(fr, q)
    prev = np.concatenate([p, O.matrix_to_quaternion(R)], 1)
    tgt = O.compute_target_pose(prev, a, True, "root_translation:root_aligned_body_rotation")
    return fr, q, a, tgt


def _timed(fn, n_items: int, budget_s: float, n0: int = 512, cap: int = 1 << 20):
    """items/s of ``fn(n)`` on a bounded sample: a short probe sizes a run of about ``budget_s`` seconds."""
    t = time.perf_counter()
    fn(n0)
    r0 = n0 / (time.perf_counter() - t)
    S = int(min(max(r0 * budget_s, n0), cap, n_items))
    t = time.perf_counter()
    out = fn(S)
    dt = time.perf_counter() - t
    return S / dt, S, dt, out


def cpu_rates(budget_s: float, cores: int, B: int):
    """The reference's CPU arithmetic for this path, restated in oracle/ik_oracle.c (fp64; pinocchio / sapien absent):
      path A port : J + pinv + one step (kinematics.py:156-172) -- the SAME algorithm the GPU arm runs (like for like);
      CLIK        : what the reference actually runs on sim_backend=cpu (kinematics.py:173-186 ->
                    PinocchioModel.compute_inverse_kinematics, <= 100 iterations) -- a different algorithm, reported beside it;
    each on all host cores (batched C loop) and single-threaded ONE ENV PER CALL through python -> ctypes, which is how
    the reference's CPU path is driven (num_envs = 1, kinematics.py:174-186)."""
    from oracle import c_oracle as CO

    CO.build()
    fr, q, a, tgt = cpu_inputs(min(B, 1 << 20))
    n = q.shape[0]
    ra, Sa, dta, _ = _timed(lambda k: CO.delta_step(fr, q[:k], a[:k], threads=cores), n, budget_s * 0.35)
    rc, Sc, dtc, oc = _timed(lambda k: CO.clik(fr, q[:k], tgt[:k], threads=cores), n, budget_s * 0.35)
    conv = float(oc[1].mean())

    def one_by_one(fn):
        def run(k):
            for e in range(k):
                fn(e)
        return run

    r1a, S1a, _, _ = _timed(one_by_one(lambda e: CO.delta_step(fr, q[e:e + 1], a[e:e + 1], threads=1)), n, budget_s * 0.1, n0=64)
    r1c, S1c, _, _ = _timed(one_by_one(lambda e: CO.clik(fr, q[e:e + 1], tgt[e:e + 1], threads=1)), n, budget_s * 0.2, n0=64)
    return {
        "path_a": (ra, Sa, dta), "clik": (rc, Sc, dtc, conv),
        "path_a_single_thread_one_env_per_call": (r1a, S1a), "clik_single_thread_one_env_per_call": (r1c, S1c),
    }


def cpu_baseline_block(r, cores: int):
    ra, Sa, dta = r["path_a"]
    rc, Sc, dtc, conv = r["clik"]
    return {
        "value": ra, "unit": "envs*steps/s", "cores": cores, "kind": "port",
        "sample": (f"{Sa} envs of the same workload in {dta:.2f} s: path A (J + pinv + one step, kinematics.py:156-172; the algorithm the GPU "
                   f"arm runs), fp64 C restatement oracle/ik_oracle.c, all {cores} host threads"),
        "algorithm": "path A (like for like with the GPU arm)",
        "single_thread_value": r["path_a_single_thread_one_env_per_call"][0],
        "single_thread_note": "one env per call through python -> ctypes -> C, one thread: how the reference drives its CPU path (num_envs = 1)",
        "clik_value": rc,
        "clik_single_thread_value": r["clik_single_thread_one_env_per_call"][0],
        "clik_note": (f"the reference's own CPU algorithm (CLIK <= 100 iterations, kinematics.py:173-186; eps / dt / damp recalled, pinocchio "
                      f"absent): {Sc} envs in {dtc:.2f} s, converged {conv:.3f}; a different algorithm from path A, NOT like for like"),
    }


def run_reference(args):
    """--impl reference: the reference's CPU arithmetic for this path on the host cores.  The headline value is the
    path-A port (the algorithm the GPU arm runs); CLIK -- what sim_backend=cpu actually runs -- rides along."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    from oracle import c_oracle as CO

    CO.build()
    cores = os.cpu_count() or 1
    B = args.envs
    fr, q, a, tgt = cpu_inputs(min(B, 1 << 20))
    n = q.shape[0]
    t = time.perf_counter()
    CO.delta_step(fr, q[:4096], a[:4096], threads=cores)
    rate0 = 4096 / (time.perf_counter() - t)
    total = max(1, args.steps + args.warmup)
    S = int(min(max(rate0 * min(120.0 / total, 4.0), 1024), n))  # envs per timed step: the whole batch when it fits
    for _ in range(args.warmup):
        CO.delta_step(fr, q[:S], a[:S], threads=cores)
    t = time.perf_counter()
    for _ in range(args.steps):
        CO.delta_step(fr, q[:S], a[:S], threads=cores)
    dt = time.perf_counter() - t
    v = S * args.steps / dt
    extra = cpu_rates(min(args.cpu_budget, 12.0), cores, B)
    cb = cpu_baseline_block(extra, cores)
    cb["value"] = v
    cb["sample"] = (f"{S} of {B} envs per step x {args.steps} steps in {dt:.2f} s (same state / action distribution): path A "
                    f"(J + pinv + one step, kinematics.py:156-172), fp64 C restatement oracle/ik_oracle.c, all {cores} host threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "envs*steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": bench_config(B),
        "envs_per_step_timed": S, "ms_per_step_full_batch_extrapolated": 1e3 * dt / args.steps * (B / S),
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": "envs*steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def bench_config(B: int) -> dict:
    """The part of `config` both arms share (the driver compares them)."""
    return {"workload": workload_name(B), "envs_per_gpu": B, "robot": "panda", "control_mode": "pd_ee_delta_pose"}


def workload_name(B: int) -> str:
    cfg = "configs[1]" if B == 4096 else "configs[4]"
    return f"{cfg}: synthetic Panda pd_ee_delta_pose (PickCube-v1 controller config), {B} envs per GPU, rest keyframe + N(0,0.02), actions U(-1,1)"


# ----------------------------------------------------------------------------- GPU arm
def gpu_ref_restated(B: int, reps: int = 30):
    """Reference GPU path at B envs: eager-torch restatement of pk chain.jacobian +
    torch.linalg.pinv + index_put scatter (baseline/eager_ref.py; pytorch_kinematics itself is
    not installable).  Returns ms per controller step (CUDA events, L2 flushed between)."""
    import torch
    from baseline.eager_ref import EagerDeltaController

    qpos = torch.tensor(PANDA_REST, dtype=torch.float32, device="cuda")[None] + 0.02 * torch.randn(B, 9, device="cuda")
    tgt = torch.zeros_like(qpos)
    idx = torch.arange(B, device="cuda")
    ctrl = EagerDeltaController(URDF, EE, qpos, tgt, idx, list(range(7)), list(range(7)))
    action = 2 * torch.rand(B, 6, device="cuda") - 1
    for _ in range(3):
        ctrl.set_action(action)
    torch.cuda.synchronize()
    flush = torch.empty(2 * L2_BYTES // 4, dtype=torch.float32, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        flush.zero_()
        a.record()
        ctrl.set_action(action)
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / reps


def pcie_floor(B: int, reps: int) -> float:
    """Seconds per step of the e2e leg's bare transfers on this rank: B x 64 B host->device and B x 28 B device->host from /
    to pinned memory on two streams, back to back, no kernel (the bound of Kinematics.delta_step_host)."""
    import torch

    h_in = torch.empty(B * 64, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(B * 28, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(B * 64, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(B * 28, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def both():
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)

    both()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        both()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def shard_statistics(group, B: int, K: int, ms: float, rank: int):
    """This rank's ShardStats with MEASURED columns: |joint target - fp64 oracle| of a 4096-env sample of the shard (path A,
    through the controller step just timed) and the iteration count of a 4096-env path-B solve (msik_target_ik iters_out)."""
    import torch
    from maniskill_b200 import sharding
    from oracle import c_oracle as CO
    from oracle import oracle as O

    CO.build()
    n = min(B, 4096)
    idx = torch.linspace(0, B - 1, n, device="cuda").long()
    group.step()
    torch.cuda.synchronize()
    q = group.art.px.cuda_articulation_qpos.torch()[idx].cpu().numpy().astype(np.float64)
    act = group.action[idx].cpu().numpy().astype(np.float64)
    got = group.art.px.cuda_articulation_target_qpos.torch()[idx, :7].cpu().numpy()
    fr = O.load_chain(URDF, EE)
    ref = CO.delta_step(fr, q[:, :7], O.pose_clip_and_scale(act[:, :6], -0.1, 0.1, -0.1), threads=4)
    err = np.abs(got - ref).max(1)
    # path B on the same sample: absolute targets = FK of a perturbed state
    kin = group.ctrl.kinematics
    qs = group.art.px.cuda_articulation_qpos.torch()[idx].contiguous()
    g = torch.Generator(device="cuda").manual_seed(7 + rank)
    q2 = qs.clone()
    q2[:, :7] += 0.3 * (torch.rand((n, 7), generator=g, device="cuda") - 0.5)
    _, info = kin.compute_ik(kin.forward_kinematics(q2), qs, return_info=True)
    iters = float(info["iterations"].abs().float().mean())
    return sharding.ShardStats(solves=float(B * K), elapsed_ms=ms, err_sum=float(err.mean()) * B * K, err_max=float(err.max()),
                               iterations=iters * B * K)


def small_batch_latency(group, reps: int = 2000):
    """Host-side cost of ONE controller step at a small env count (buffers L2-resident, steps back to back):
    eager issue through the public API vs the same step captured once in a CUDA graph and replayed."""
    import torch

    for _ in range(20):
        group.step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        group.step()
    torch.cuda.synchronize()
    eager = (time.perf_counter() - t0) / reps
    ctrl = group.combined or group.ctrl
    out = {"us_per_step_eager": 1e6 * eager}
    if hasattr(ctrl, "capture"):
        replay = ctrl.capture(group.action)
        for _ in range(20):
            replay()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            replay()
        torch.cuda.synchronize()
        out["us_per_step_graph"] = 1e6 * (time.perf_counter() - t0) / reps
    # a fresh action tensor every step (what a policy hands over): the layout checks run every time
    acts = [group.action.clone() for _ in range(4)]
    step_fn = ctrl.set_action if group.combined is not None else (lambda a: group.ctrl.set_action_flat(a, 0, group.adim))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(reps):
        step_fn(acts[i & 3])
    torch.cuda.synchronize()
    out["us_per_step_eager_new_action_tensor"] = 1e6 * (time.perf_counter() - t0) / reps
    # 16 steps per graph replay: the launch-issue floor amortised (device time per step is what is left)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(16):
            group.step()
    gr.replay()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps // 16):
        gr.replay()
    torch.cuda.synchronize()
    out["us_per_step_graph_of_16_steps"] = 1e6 * (time.perf_counter() - t0) / (reps // 16) / 16
    out["latency_note"] = ("a step is ONE kernel: eager = prepared C-ABI step (msik_step_launch) behind CombinedController.set_action; a one-node "
                           "graph replay costs what one launch costs (cudaGraphLaunch + torch's replay bookkeeping); graphs pay only "
                           "when several steps are replayed together")
    return out


def launch_floor_us(reps: int = 50) -> float:
    """Device time of the smallest possible launch under the small-batch timing recipe (L2 flushed, one event pair per step)."""
    import torch

    x = torch.zeros(1, device="cuda")
    flush = torch.empty(2 * L2_BYTES // 4, dtype=torch.float32, device="cuda")
    for _ in range(3):
        x.add_(1.0)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        flush.zero_()
        a.record()
        x.add_(1.0)
        b.record()
    torch.cuda.synchronize()
    return 1e3 * sum(a.elapsed_time(b) for a, b in ev) / reps


def graph_hot_us(group, n: int = 16, reps: int = 50) -> float:
    """Device time per step of n back-to-back steps replayed from one CUDA graph on L2-resident buffers."""
    import torch

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            group.step()
    torch.cuda.current_stream().wait_stream(side)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(n):
            group.step()
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / (reps * n)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from maniskill_b200 import _lib as L
    from maniskill_b200 import sharding

    rank, local_rank, world = sharding.env_world()
    global SEED_BASE
    SEED_BASE = 64 * rank
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the product has no CPU path"
    torch.cuda.set_device(local_rank)
    if world > 1:
        sharding.init_distributed("nccl")
    lib = L.load()
    L.require_device()
    B, K, W = args.envs, args.steps, args.warmup
    peak, peak_src = peaks()
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    try:
        phys = int(vis.split(",")[local_rank]) if vis else local_rank
    except ValueError:
        phys = local_rank
    sampler = ClockSampler(phys)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    groups, l2mode = make_groups("panda", B)
    n0 = lib.msik_launch_count()
    groups[0].step()
    launches_per_step = lib.msik_launch_count() - n0

    # (1) device-resident throughput through the public controller API, eager issue
    barrier()
    ms_eager = max_over_ranks(time_steps(groups, l2mode, K, W, sampler))
    # (2) same K steps as one CUDA graph (ring mode only): this is `value`
    if l2mode == "ring" and K <= 8192:
        barrier()
        ms = max_over_ranks(time_graph(groups, K, W, sampler))
        launch = "cuda_graph"
    else:
        ms, launch = ms_eager, "eager"
    if ms_eager < ms:
        ms, launch = ms_eager, "eager"
    value = world * B * K / (ms * 1e-3)
    kernel_ms = ms / K
    bps = BYTES_PER_SOLVE["panda_pose"]
    achieved = bps * B / (kernel_ms * 1e-3) / 1e9

    # (3) end to end with HOST buffers through Kinematics.delta_step_host (C ABI msik_delta_step_host)
    kin = groups[0].ctrl.kinematics
    h_q = groups[0].art.px.cuda_articulation_qpos.torch().cpu().pin_memory()
    h_a = groups[0].action.cpu().pin_memory()
    h_o = torch.empty((B, 7), dtype=torch.float32).pin_memory()
    norm = ((-0.1,) * 3, (0.1,) * 3, -0.1)
    Ke = max(3, min(K, int(2e9 // max(B * 92, 1))))  # bound host traffic of the e2e leg
    for _ in range(3):
        kin.delta_step_host(h_q, h_a, h_o, normalize=norm, chunk_envs=args.chunk)
    barrier()
    sampler.active = True
    t0 = time.perf_counter()
    for _ in range(Ke):
        kin.delta_step_host(h_q, h_a, h_o, normalize=norm, chunk_envs=args.chunk)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    sampler.active = False
    e2e = world * B * Ke / e2e_s
    # the host result is the same joint targets the resident path wrote
    ref_rows = groups[0].art.px.cuda_articulation_target_qpos.torch()[:, :7].cpu()
    e2e_maxdiff = float((h_o - ref_rows).abs().max())

    # PCIe floor of the e2e leg on THIS rank, all ranks copying at once (the box's host memory / PCIe root is shared):
    # the same bytes per step, one host->device and one device->host stream, no kernel
    barrier()
    floor_s = max_over_ranks(pcie_floor(B, max(3, min(Ke, 10))))
    # per-shard statistics of this rank's rollout (the ONLY collective of the job, after the timed regions):
    # path A joint targets of a 4096-env sample against the fp64 C oracle, and a 4096-env path-B solve
    # (absolute targets) for the iteration / residual columns
    stats = shard_statistics(groups[0], B, K, ms, rank)
    allst = sharding.gather_stats(stats, device="cuda")
    clocks = sampler.summary()
    if rank != 0:
        sharding.shutdown_distributed()
        return 0

    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            traffic = json.load(f).get(str(B))
    line = {
        "metric": METRIC, "value": value, "unit": "envs*steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {**bench_config(B),
                   "kernel": kin.kernel_name, "launch": launch,
                   "step": ("CombinedController(arm pd_ee_delta_pose + mimic gripper).set_action, one fused launch writing whole "
                            "target rows" if FUSED else "PDEEPoseController.set_action (arm only; old target rows are read to "
                            "preserve the finger columns)"), "l2": ("ring of %d env groups (%.0f MB > L2)" % (
                       len(groups), len(groups) * groups[0].set_bytes() / 1e6)) if l2mode == "ring" else "L2 flushed between steps",
                   "parallelism": f"env-sharded x{world}, no collective in the step"},
        "value_eager_api": world * B * K / (ms_eager * 1e-3),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "bytes_per_env": bps, "kernel_us": kernel_ms * 1e3, "peak_source": peak_src},
        "e2e": {"value": e2e, "unit": "envs*steps/s", "h2d_bytes_per_step": B * (9 + 7) * 4, "d2h_bytes_per_step": B * 7 * 4,
                "steps": Ke, "ms_per_step": 1e3 * e2e_s / Ke, "api": "Kinematics.delta_step_host -> msik_delta_step_host (pinned host buffers)",
                "max_abs_diff_vs_resident": e2e_maxdiff, "pcie_floor_ms_per_step": 1e3 * floor_s,
                "frac_of_pcie_floor": floor_s / (e2e_s / Ke),
                "pcie_floor_note": "same bytes per step copied host->device and device->host on two streams by all ranks at once, no kernel"},
        "gpu_launches": int(K * launches_per_step),
        "clocks": clocks,
        "shard_stats": {**sharding.reduce_stats(allst),
                        "per_rank": [{f: getattr(st, f) for f in sharding.STAT_FIELDS} for st in allst],
                        "note": "err_* = |joint target - fp64 oracle| (rad) on a 4096-env sample of each rank's shard (path A); iterations "
                                "= solver iterations of a 4096-env path-B solve on that rank (msik_target_ik iters_out)"},
    }
    if world == 1 and not args.no_extras:
        cores = os.cpu_count() or 1
        line["cpu_baseline"] = cpu_baseline_block(cpu_rates(args.cpu_budget, cores, B), cores)
        # configs[1]: the RL-typical 4096 envs, where the host issue path is the step time
        g4, m4 = make_groups("panda", 4096)
        ms4 = time_steps(g4, m4, 50, 5) / 50
        lat = small_batch_latency(g4[0])
        ref_ms = gpu_ref_restated(4096)
        line["gpu_reference_4096"] = {"ours_ms": ms4, "reference_restated_ms": ref_ms, "speedup": ref_ms / ms4, **lat,
                                      "note": "reference GPU path restated in eager torch (pytorch_kinematics 0.7.5 not installable); "
                                              "ours_ms: L2 flushed between steps; us_per_step_*: back-to-back steps on L2-resident buffers "
                                              "(host issue cost vs CUDA-graph replay)"}
        del g4
        # the other BASELINE.json configs, timed the same way (they are parity-test cases; reported for context)
        others = {}
        gs, m = make_groups("panda", 65536, "pd_ee_pose")
        k = 20
        t_ms = time_steps(gs, m, k, 3)
        iters = torch.zeros(65536, dtype=torch.int32, device="cuda")
        c0 = gs[0].ctrl
        import ctypes as C
        qpos0, tgt0, _ = c0._physx()
        cfg0 = c0._step_cfg(gs[0].action.stride(0), 0, c0._flags())
        it0 = c0.kinematics.iter_cfg()
        L.check(lib.msik_target_ik(c0.kinematics._chain, C.byref(cfg0), C.byref(it0), L.ptr(qpos0), None, L.ptr(gs[0].action), None,
                                   L.ptr(c0._dense("_tp_buf", 7)), L.ptr(tgt0), None, None, L.ptr(iters), None,
                                   torch.cuda.current_stream().cuda_stream))
        us2 = 1e3 * t_ms / k
        mean_it = float(iters.abs().float().mean())
        flop = 65536 * mean_it * IK_FLOP_PER_ENV_ITER
        others["configs[2] Panda pd_ee_pose absolute (iterative IK), 65536 envs"] = {
            "us_per_step": us2, "envs_steps_per_s": 65536 * k / (t_ms * 1e-3), "mean_iterations": mean_it,
            "max_iterations": int(iters.abs().max()), "envs_on_damped_step": int((iters < 0).sum()), "l2": m,
            "roofline": {"bound": "fp32 issue (HBM is irrelevant: 80 B per env per ~22 iterations)",
                         "achieved": flop / (us2 * 1e-6) / 1e12, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s",
                         "frac": flop / (us2 * 1e-6) / 1e12 / FP32_PEAK_TFLOPS,
                         "flop_per_env_iteration": IK_FLOP_PER_ENV_ITER,
                         "note": "useful fp32 flop = envs x mean iterations x the loop body's SASS count (FFMA = 2; profiles/r02*_target_ik_*), "
                                 "peak = 148 SMs x 128 lanes x 2 x max SM clock; issue-slot utilisation is in the ncu summary of the same kernel"}}
        del gs
        gs, m = make_groups("fetch", 16384)
        t_ms = time_steps(gs, m, 50, 3)
        floor_us = launch_floor_us()
        hot_us = graph_hot_us(gs[0])
        others["configs[3] Fetch pd_ee_delta_pose (7 of 11 ancestors), 16384 envs"] = {
            "us_per_step": 1e3 * t_ms / 50, "envs_steps_per_s": 16384 * 50 / (t_ms * 1e-3), "kernel": gs[0].ctrl.kinematics.kernel_name,
            "l2": m,
            "bound": {"kind": "launch + first-touch DRAM latency (16384 envs = 512 warps = 0.9 per SM sub-partition: one dependent pass "
                              "load -> solve -> store, nothing to overlap it with); HBM is at 3 %",
                      "smallest_launch_us_same_timing": floor_us, "frac_of_floor": floor_us / (1e3 * t_ms / 50),
                      "us_per_step_l2_hot_graph_of_16": hot_us,
                      "note": "smallest_launch = a 1-element torch kernel timed the same way (L2 flushed, event pair per step): what ANY kernel "
                              "costs in this measurement; l2_hot = 16 back-to-back steps per CUDA-graph replay on L2-resident buffers"}}
        line["gpu_reference_4096"]["bound"] = {
            "kind": "launch + first-touch DRAM latency (4096 envs = 128 warps on 592 sub-partitions)",
            "smallest_launch_us_same_timing": floor_us, "frac_of_floor": floor_us / (1e3 * line["gpu_reference_4096"]["ours_ms"])}
        del gs
        torch.cuda.empty_cache()
        # configs[0]: one env, as the reference's CPU backend steps it (there: Pinocchio CLIK on the host; here the same controller
        # on the GPU -- the product has no CPU path): wall time of set_action + synchronize, i.e. what a num_envs = 1 loop sees
        g1 = Group("panda", 1, SEED_BASE, "pd_ee_delta_pose", True)
        for _ in range(50):
            g1.step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(2000):
            g1.step()
            torch.cuda.synchronize()
        us1 = 1e6 * (time.perf_counter() - t0) / 2000
        cb = line["cpu_baseline"]
        others["configs[0] PickCube Panda pd_ee_delta_pose, num_envs=1 (GPU controller; the reference runs Pinocchio CLIK on the CPU here)"] = {
            "us_per_step_wall_with_sync": us1, "steps_per_s": 1e6 / us1,
            "cpu_single_thread_clik_us_per_step": 1e6 / cb["clik_single_thread_value"],
            "cpu_single_thread_path_a_us_per_step": 1e6 / cb["single_thread_value"],
            "note": "CPU figures: oracle/ik_oracle.c, one env per call through python -> ctypes (cpu_baseline.*single_thread*)"}
        del g1
        # the fork's own robot (not a BASELINE config; reported because Zhouxunzhe/ManiSkill's addition IS this robot): Realman,
        # stock pd_ee_delta_pos dict = 2 arms + 2 six-finger grippers + head + base, ONE launch (msik_delta_step_dual)
        from maniskill_b200 import presets

        for b in (4096, 1 << 20):
            n0 = lib.msik_launch_count()
            gs = [PresetGroup(presets.realman_pd_ee_delta_pos(), presets.REALMAN_URDF, ("r_link6", "l_link6"), b, SEED_BASE + i)
                  for i in range(3 if b > 100000 else 1)]
            gs[0].step()
            n1 = lib.msik_launch_count()
            gs[0].step()
            nl = lib.msik_launch_count() - n1
            k = 30
            t_ms = time_graph(gs, k, 3) if len(gs) > 1 else graph_hot_us(gs[0]) * k * 1e-3
            moved = (41 * 3 + 22) * 4  # qpos row + old target row + new target row + action row (row-granular)
            others[f"fork robot: Realman pd_ee_delta_pos stock dict (mobile_realman.py:436-443), {b} envs"] = {
                "us_per_step": 1e3 * t_ms / k, "envs_steps_per_s": b * k / (t_ms * 1e-3), "launches_per_step": int(nl),
                "row_bytes_moved_per_env": moved, "GBps_moved": moved * b / (t_ms / k * 1e-3) / 1e9,
                "l2": "ring of 3 env groups, CUDA graph" if len(gs) > 1 else "L2-hot, 16 steps per CUDA-graph replay"}
            del gs
            torch.cuda.empty_cache()
        line["other_configs"] = others
        if not args.no_sweep:
            sweep = []
            for b in (4096, 16384, 65536, 262144, 1 << 20, 1 << 22):
                gs, m = make_groups("panda", b)
                k = 64
                t_ms = time_graph(gs, k, 3) if m == "ring" else time_steps(gs, m, k, 3)
                gbs = bps * b / (t_ms / k * 1e-3) / 1e9
                sweep.append({"envs": b, "us_per_step": 1e3 * t_ms / k, "envs_steps_per_s": b * k / (t_ms * 1e-3), "GBps": gbs,
                              "frac": gbs / peak, "l2": m})
                del gs
                torch.cuda.empty_cache()
            line["sweep"] = sweep
    emit(line)
    sharding.shutdown_distributed()
    return 0


_REAL_STDOUT = None


def _own_stdout():
    """bench.py's stdout carries exactly ONE JSON line: move fd 1 to stderr for everybody else
    (NCCL prints its version banner to stdout) and keep a private handle on the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line: dict):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    _own_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--envs", type=int, default=1 << 20, help="envs per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--chunk", type=int, default=1 << 18, help="envs per chunk of the host pipeline")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for cpu_baseline")
    ap.add_argument("--no-extras", action="store_true", help="skip cpu_baseline / GPU reference / sweep")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--arm-only", action="store_true",
                    help="step = PDEEPoseController.set_action alone (default: the stock Panda CombinedController, arm "
                         "pd_ee_delta_pose + mimic gripper, whose set_action is ONE fused launch)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world != args.gpus:
        args.gpus = world
    global FUSED
    FUSED = not args.arm_only
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
