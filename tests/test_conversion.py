"""Trajectory action conversion row: the batched pose -> pd_ee_delta action mapping
(conversion.py:52-81,152-183) against the oracle, and as a round trip through the controller."""
import numpy as np
import pytest
import torch
from scipy.spatial.transform import Rotation

from conftest import PANDA
from oracle import oracle as O


def _poses(B, seed, max_step=0.08, max_rot=0.09):
    rng = np.random.default_rng(seed)
    ee = np.concatenate([rng.normal(0, 0.5, (B, 3)), O.standardize_quaternion(Rotation.random(B, random_state=seed).as_quat()[:, [3, 0, 1, 2]])], 1)
    rv = rng.normal(size=(B, 3))
    rv *= (rng.uniform(0, max_rot, (B, 1)) / np.linalg.norm(rv, axis=1, keepdims=True))
    dq = Rotation.from_rotvec(rv).as_quat()[:, [3, 0, 1, 2]]
    tq = O.quaternion_raw_multiply(dq, ee[:, 3:])  # target = delta (x) ee  (root-aligned rotation)
    tgt = np.concatenate([ee[:, :3] + rng.uniform(-max_step, max_step, (B, 3)), tq], 1)
    return tgt, ee, rv


def test_oracle_axis_angle_against_scipy_and_round_trip():
    tgt, ee, rv = _poses(512, 1)
    a, exact = O.pose_to_delta_action(tgt, ee, -0.1, 0.1)
    assert exact.all()
    # delta_q = ee (x) target^-1 = (delta)^-1: the compact axis-angle is -rotvec; / 0.1 by inv_scale_action
    assert np.abs(a[:, 3:] * 0.1 + rv).max() < 1e-12
    assert np.abs(a[:, :3] * 0.1 - (tgt[:, :3] - ee[:, :3])).max() < 1e-15
    # the controller multiplies the rotation part by rot_lower = -0.1 (pd_ee_pose.py:209): the sign comes back
    proc = O.pose_clip_and_scale(a, -0.1, 0.1, -0.1)
    assert np.abs(proc[:, 3:] - rv).max() < 1e-12
    # and its euler-XYZ target agrees with the axis-angle one to second order in the step
    back = O.compute_target_pose(ee, proc, True, "root_translation:root_aligned_body_rotation")
    assert np.abs(back[:, :3] - tgt[:, :3]).max() < 1e-12
    d = np.minimum(np.abs(back[:, 3:] - O.standardize_quaternion(tgt[:, 3:])).max(1),
                   np.abs(back[:, 3:] + O.standardize_quaternion(tgt[:, 3:])).max(1))
    assert d.max() < 0.09 ** 2


def test_oracle_clipping_and_identity_rules():
    tgt, ee, _ = _poses(256, 2, max_step=0.3, max_rot=0.4)
    a, exact = O.pose_to_delta_action(tgt, ee, -0.1, 0.1)
    assert (~exact).any() and exact.any()
    assert np.abs(a[:, :3]).max() <= 1 and np.linalg.norm(a[:, 3:], axis=1).max() <= 1 + 1e-12
    same, _ = O.pose_to_delta_action(ee, ee, -0.1, 0.1)  # no motion: identity rotation -> zero action
    assert np.abs(same).max() == 0
    a3, e3 = O.pose_to_delta_action(tgt, ee, -0.1, 0.1, pos_only=True)
    assert a3.shape == (256, 3) and np.array_equal(a3, a[:, :3])


@pytest.mark.gpu
@pytest.mark.parametrize("B,step,rot", [(4096, 0.08, 0.09), (1000, 0.3, 0.4), (1, 0.05, 3.0), (65536, 0.08, 0.09)])
def test_gpu_pose_to_delta_action_matches_oracle(B, step, rot):
    from helpers import make_art, pose_cfg
    from maniskill_b200 import trajectory as TR
    from maniskill_b200.structs.pose import Pose

    art = make_art(PANDA, B)
    cfg = pose_cfg(PANDA)
    ctrl = cfg.controller_cls(cfg, art, control_freq=20, sim_freq=100, scene=art.scene)
    tgt, ee, _ = _poses(B, 3, step, rot)
    tgt32, ee32 = tgt.astype(np.float32), ee.astype(np.float32)
    act, exact = TR.delta_pose_to_pd_ee_delta(ctrl, Pose.create(torch.from_numpy(tgt32).cuda()), Pose.create(torch.from_numpy(ee32).cuda()))
    ref, rex = O.pose_to_delta_action(tgt32.astype(np.float64), ee32.astype(np.float64), -0.1, 0.1)
    err = np.abs(act.cpu().numpy() - ref).max(1)
    # fp32 quaternions carry ~1e-7 of angle noise, / 0.1 by the inverse scaling
    assert err.max() < 2e-5
    border = np.abs(np.abs(ref[:, :3]).max(1) - 1) < 1e-5  # clip decisions may flip exactly at the bound
    assert (exact.cpu().numpy() == rex)[~border].mean() > 0.999
    a3, _ = TR.delta_pose_to_pd_ee_delta(ctrl, Pose.create(torch.from_numpy(tgt32).cuda()), Pose.create(torch.from_numpy(ee32).cuda()),
                                         pos_only=True)
    assert torch.equal(a3, act[:, :3])


@pytest.mark.gpu
def test_gpu_conversion_round_trip_through_the_controller():
    """qpos demonstration step -> target EE pose (FK kernel) -> pd_ee_delta_pose action (conversion kernel)
    -> controller.set_action (path A) -> FK of the new joint targets lands on the target pose
    (first order: one Jacobian step, euler vs axis-angle second-order difference)."""
    from helpers import PANDA_REST, fill_qpos, make_art, pose_cfg, sample_q, set_link_poses_from_fk
    from maniskill_b200 import trajectory as TR

    B = 8192
    art = make_art(PANDA, B)
    cfg = pose_cfg(PANDA)
    ctrl = cfg.controller_cls(cfg, art, control_freq=20, sim_freq=100, scene=art.scene)
    fr = O.load_chain(PANDA[0], PANDA[1])
    fill_qpos(art, fr, list(range(7)), sample_q(fr, B, seed=7, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    q = art.get_qpos()
    demo_target = q[:, :7] + (torch.rand(B, 7, device="cuda") - 0.5) * 0.06  # the pd_joint_pos demo's next target
    tgt_pose = TR.target_ee_pose_from_qpos(ctrl, q, demo_target)
    act, exact = TR.delta_pose_to_pd_ee_delta(ctrl, tgt_pose)
    assert exact.float().mean().item() > 0.99
    ctrl.set_action(act)
    new_q = q.clone()
    new_q[:, :7] = art.drive_targets[:, :7]
    reached = art.pose * ctrl.kinematics.forward_kinematics(new_q)
    perr = (reached.p - tgt_pose.p).norm(dim=1)[exact]
    moved = (tgt_pose.p - ctrl.ee_link.pose.p).norm(dim=1)[exact]
    assert (perr < 0.05 * moved + 1e-4).float().mean().item() > 0.99  # one linearised step closes >95 % of the gap


@pytest.mark.gpu
@pytest.mark.parametrize("robot_name", ["panda", "fetch"])
def test_gpu_world_frame_fk_matches_oracle(robot_name):
    """tcp / EE world pose from qpos alone (root pose * FK in one launch) against the fp64
    oracle, 1e-6 m / 1e-6 quaternion."""
    from conftest import FETCH
    from helpers import make_art, sample_q
    from maniskill_b200.kinematics import Kinematics
    from maniskill_b200.structs.pose import Pose

    robot = PANDA if robot_name == "panda" else FETCH
    B = 4099
    art = make_art(robot, B)
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in robot[2]], dtype=torch.int32, device="cuda")
    kin = Kinematics(robot[0], robot[1], art, idx)
    fr = O.load_chain(robot[0], robot[1])
    anc = kin.active_ancestor_joint_idxs
    anc = anc.cpu().numpy() if hasattr(anc, "cpu") else np.asarray(anc)
    qa = sample_q(fr, B, seed=11)
    if robot_name == "fetch":
        qa[:, :2] = np.random.default_rng(5).uniform(-2, 2, (B, 2))  # keep the planar base near the origin for the 1e-6 m bar
    rows = np.zeros((B, art.max_dof), np.float32)
    rows[:, anc] = qa
    qa = rows[:, anc].astype(np.float64)
    rng = np.random.default_rng(12)
    root = np.concatenate([rng.uniform(-1, 1, (B, 3)), Rotation.random(B, random_state=13).as_quat()[:, [3, 0, 1, 2]]], 1).astype(np.float32)
    got = kin.forward_kinematics(torch.from_numpy(rows).cuda(), root_pose=Pose.create(torch.from_numpy(root).cuda())).raw_pose.cpu().numpy()
    _, R, p = O.jacobian(fr, qa)
    ref = O.pose_mul(root.astype(np.float64), np.concatenate([p, O.matrix_to_quaternion(R)], 1))
    assert np.abs(got[:, :3] - ref[:, :3]).max() < (1e-6 if robot_name == "panda" else 2e-6)
    dq = np.minimum(np.abs(got[:, 3:] - ref[:, 3:]).max(1), np.abs(got[:, 3:] + ref[:, 3:]).max(1))
    assert dq.max() < 1e-6
    # same numbers as the two-step form (FK kernel, then Pose.__mul__ in torch) to fp32 rounding
    two = (Pose.create(torch.from_numpy(root).cuda()) * kin.forward_kinematics(torch.from_numpy(rows).cuda())).raw_pose.cpu().numpy()
    assert np.abs(two[:, :3] - got[:, :3]).max() < 1e-6
