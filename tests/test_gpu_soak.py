"""Bitwise-repeat soak of EVERY refill path of the persistent TMA kernel.

compute-sanitizer is not usable on this pool, and the one race the kernel ever had lived in a refill ordering (a bulk
refill of a shared-memory stage overtaking late reads of the previous tile, msik_tma.cuh).  Each instantiation below
takes a different refill / store path; each is launched repeatedly on L2-hot inputs (the condition under which the race
showed) and every result must equal the first one bit for bit:

  lean            stock Panda dict (arm + mimic gripper): LEAN = 1 instantiation, whole-row bulk stores, old targets never read
  pos_lean        the same dict with pd_ee_delta_pos: the 3-dof LEAN = 1 instantiation (different geometry of the action tile)
  arm_only        PDEEPoseController alone: LEAN = 2, load_t path (old target rows re-read through the same mbarrier, merged, stored)
  eager_state     state_mode = "eager": non-LEAN instantiation with per-lane tq / sq stores next to the bulk stores
  pos_only        pd_ee_delta_pos alone: 3-dof LEAN = 2 instantiation with the load_t path
  fetch_fused_vel stock Fetch dict (arm + gripper + body + base velocity): non-LEAN, per-lane qvel stores, 15-float rows
  no_drive        Kinematics.compute_ik(use_delta_ik_solver=True): NO_DRIVE instantiation, no target tile at all
"""
import numpy as np
import pytest
import torch

from conftest import FETCH, PANDA
from helpers import PANDA_REST, make_art, pos_cfg, pose_cfg, set_link_poses_from_fk
from maniskill_b200.controllers import (
    CombinedController,
    PDBaseVelControllerConfig,
    PDJointPosControllerConfig,
    PDJointPosMimicControllerConfig,
)

pytestmark = pytest.mark.gpu
B = (1 << 18) + 37  # 8193 warp tiles over 2960 resident warps: every warp refills its stages, plus a ragged tail
REPS = 40
FETCH_REST = [0, 0, 0, 0.386, 0, 0, 0, -np.pi / 4, 0, np.pi / 4, 0, np.pi / 3, 0, 0.015, 0.015]
GRIP = dict(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01, upper=0.04, stiffness=1e3, damping=1e2, force_limit=100)


def _fill(art, rest):
    g = torch.Generator(device="cuda").manual_seed(99)
    q = art.px.cuda_articulation_qpos.torch()
    q.copy_(torch.tensor(rest, dtype=torch.float32, device="cuda")[None] + 0.02 * torch.randn(q.shape, generator=g, device="cuda"))
    return g


def _soak(step, outputs):
    step()
    ref = [o.clone() for o in outputs()]
    for r in range(REPS):
        step()
        for a, b in zip(outputs(), ref):
            assert torch.equal(a, b), f"repeat {r}: {(a != b).any(dim=-1).sum().item()} rows differ"


def _panda(kind):
    art = make_art(PANDA, B)
    g = _fill(art, PANDA_REST + [0.04, 0.04])
    tgt = art.px.cuda_articulation_target_qpos.torch()
    if kind in ("lean", "pos_lean"):
        arm_cfg = pose_cfg(PANDA) if kind == "lean" else pos_cfg(PANDA)
        cc = CombinedController(dict(arm=arm_cfg, gripper=PDJointPosMimicControllerConfig(**GRIP)), art, control_freq=20,
                                sim_freq=100, scene=art.scene)
        arm, adim = cc.controllers["arm"], 7 if kind == "lean" else 4
    else:
        cfg = pos_cfg(PANDA) if kind == "pos_only" else pose_cfg(PANDA)
        cc = arm = cfg.controller_cls(cfg, art, control_freq=20, sim_freq=100, scene=art.scene)
        adim = 3 if kind == "pos_only" else 6
        if kind == "eager_state":
            arm.state_mode = "eager"
    set_link_poses_from_fk(art, arm)
    cc.reset()
    action = torch.rand((B, adim), generator=g, device="cuda") * 2 - 1
    tgt.fill_(0.25)
    outs = (lambda: [tgt, arm._tq_buf, arm._sq_buf]) if kind == "eager_state" else (lambda: [tgt])
    return (lambda: cc.set_action(action)), outs


@pytest.mark.parametrize("kind", ["lean", "pos_lean", "arm_only", "eager_state", "pos_only"])
def test_panda_refill_paths_repeat_bitwise(kind):
    step, outs = _panda(kind)
    _soak(step, outs)


def test_fetch_fused_velocity_dict_repeats_bitwise():
    from maniskill_b200.shim import FakeArticulation

    art = FakeArticulation(FETCH[0], B, ee_link_names=[FETCH[1]])
    g = _fill(art, FETCH_REST)
    cfgs = dict(
        arm=pose_cfg(FETCH),
        gripper=PDJointPosMimicControllerConfig(joint_names=["l_gripper_finger_joint", "r_gripper_finger_joint"], lower=-0.01, upper=0.05,
                                                stiffness=1e3, damping=1e2, force_limit=100),
        body=PDJointPosControllerConfig(joint_names=["head_pan_joint", "head_tilt_joint", "torso_lift_joint"], lower=-0.1, upper=0.1,
                                        stiffness=1e3, damping=1e2, force_limit=100, use_delta=True),
        base=PDBaseVelControllerConfig(joint_names=["root_x_axis_joint", "root_y_axis_joint", "root_z_rotation_joint"],
                                       lower=[-1, -1, -3.14], upper=[1, 1, 3.14], damping=1000, force_limit=500),
    )
    cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
    set_link_poses_from_fk(art, cc.controllers["arm"])
    cc.reset()
    action = torch.rand((B, 13), generator=g, device="cuda") * 2 - 1
    tgt = art.px.cuda_articulation_target_qpos.torch()
    qvel = art.px.cuda_articulation_target_qvel.torch()
    tgt.fill_(0.25)
    from maniskill_b200 import _lib as L

    n0 = L.load().msik_launch_count()
    cc.set_action(action)
    assert L.load().msik_launch_count() - n0 == 1
    _soak(lambda: cc.set_action(action), lambda: [tgt, qvel])


def test_no_drive_compute_ik_repeats_bitwise():
    from maniskill_b200.kinematics import Kinematics

    art = make_art(PANDA, B)
    g = _fill(art, PANDA_REST + [0.04, 0.04])
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in PANDA[2]], dtype=torch.int32, device="cuda")
    kin = Kinematics(PANDA[0], PANDA[1], art, idx)
    q = art.px.cuda_articulation_qpos.torch()
    a = (torch.rand((B, 6), generator=g, device="cuda") * 2 - 1) * 0.1
    ref = kin.compute_ik(None, q, action=a, use_delta_ik_solver=True)
    for r in range(REPS):
        assert torch.equal(kin.compute_ik(None, q, action=a, use_delta_ik_solver=True), ref), r


def test_realman_dual_launch_repeats_bitwise():
    """Both Realman arms + grippers + head + base in ONE launch (second chain inside the persistent kernel, 2 warps per CTA
    for the 41-float rows, old target rows re-read): another refill / store geometry, same bitwise-repeat requirement."""
    from conftest import REALMAN_L, REALMAN_R
    from maniskill_b200 import _lib as L
    from maniskill_b200 import presets
    from maniskill_b200.shim import FakeArticulation
    from maniskill_b200.structs.pose import Pose

    Bn = (1 << 17) + 37
    art = FakeArticulation(REALMAN_R[0], Bn, ee_link_names=[REALMAN_R[1], REALMAN_L[1]])
    cc = CombinedController(presets.realman_pd_ee_delta_pos(REALMAN_R[0]), art, control_freq=20, sim_freq=100, scene=art.scene)
    g = torch.Generator(device="cuda").manual_seed(7)
    q = art.px.cuda_articulation_qpos.torch()
    q.copy_(1.5 * (2 * torch.rand(q.shape, generator=g, device="cuda") - 1))
    root = Pose.create(torch.tensor([[0.0, 0, 0, 1, 0, 0, 0]], device="cuda").repeat(Bn, 1))
    art.set_root_pose(root)
    for u in ("right_arm", "left_arm"):
        c = cc.controllers[u]
        art.set_link_pose(c.ee_link.name, root * c.kinematics.forward_kinematics(art.get_qpos()))
    cc.reset()
    action = torch.rand((Bn, cc.single_action_space.shape[0]), generator=g, device="cuda") * 2 - 1
    tgt = art.px.cuda_articulation_target_qpos.torch()
    qvel = art.px.cuda_articulation_target_qvel.torch()
    tgt.fill_(0.25)
    n0 = L.load().msik_launch_count()
    cc.set_action(action)
    assert L.load().msik_launch_count() - n0 == 1
    _soak(lambda: cc.set_action(action), lambda: [tgt, qvel])
