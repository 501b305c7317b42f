"""Controller-level GPU parity (PDEEPoseController / PDEEPosController on the PhysX-layout
shim) against the oracle's restatement of set_action (pd_ee_pose.py:91-115)."""
import numpy as np
import pytest
import torch

from conftest import FETCH, PANDA
from helpers import PANDA_REST, fill_qpos, make_art, pos_cfg, pose_cfg, sample_q, set_link_poses_from_fk
from maniskill_b200.controllers import (
    CombinedController,
    PDEEPoseController,
    PDJointPosMimicControllerConfig,
)
from maniskill_b200.structs.pose import Pose
from oracle import oracle as O

pytestmark = pytest.mark.gpu
Q_TOL = 1e-5


def controller(robot, cfg, B, **artkw):
    art = make_art(robot, B, **artkw)
    ctrl = cfg.controller_cls(cfg, art, control_freq=20, sim_freq=100, scene=art.scene)
    fr = O.load_chain(robot[0], robot[1])
    return art, ctrl, fr


def oracle_kw(ctrl):
    k = ctrl.kinematics
    return dict(anc_idx=k.active_ancestor_joint_idxs, ctrl_in_anc=k.controlled_joints_idx_in_qmask)


def well(fr, qa, mask=None, lim=50):
    J, _, _ = O.jacobian(fr, qa)
    if mask is not None:
        J = J[:, :, mask]
    return np.linalg.cond(J) < lim


@pytest.mark.parametrize("mode", ["lazy", "eager"])
@pytest.mark.parametrize("artkw", [dict(), dict(n_pad_rows=3, max_dof=12), dict(interleave_rows=True)])
def test_pd_ee_delta_pose_panda(mode, artkw):
    B = 1000  # ragged: not a multiple of the CTA size
    art, ctrl, fr = controller(PANDA, pose_cfg(PANDA), B, **artkw)
    ctrl.state_mode = mode
    assert ctrl.action_space.shape == (B, 6) and ctrl.single_action_space.low.min() == -1
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=1, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    tgt = art.px.cuda_articulation_target_qpos.torch()
    tgt.fill_(-7.0)
    before = tgt.clone()
    action = torch.rand(B, 6, device="cuda") * 3 - 1.5  # exercises the clip and the norm clip
    ctrl.set_action(action)
    ref = O.controller_step(fr, rows, action.cpu().numpy(), ee_pose_base=ctrl.ee_pose_at_base.raw_pose.cpu().numpy().astype(np.float64),
                            **oracle_kw(ctrl))
    got = art.drive_targets[:, :7].cpu().numpy()
    assert np.abs(got - ref["target_qpos"]).max() < Q_TOL
    # only the controlled columns of our rows changed
    after = tgt.clone()
    rowsel = art._data_index.long()
    after[rowsel, :7] = before[rowsel, :7]
    assert torch.equal(after, before)
    # reference-visible state
    assert np.abs(ctrl._target_qpos.cpu().numpy() - ref["target_qpos"]).max() < Q_TOL
    if mode == "eager":
        assert np.abs(ctrl._start_qpos.cpu().numpy() - rows[:, :7]).max() == 0
        tp = ctrl._target_pose.raw_pose.cpu().numpy()
        assert np.abs(tp[:, :3] - ref["target_pose"][:, :3]).max() < 1e-6
        qd = np.minimum(np.abs(tp[:, 3:] - ref["target_pose"][:, 3:]).max(1), np.abs(tp[:, 3:] + ref["target_pose"][:, 3:]).max(1))
        assert qd.max() < 1e-6


def test_pd_ee_delta_pos_panda():
    B = 777
    art, ctrl, fr = controller(PANDA, pos_cfg(PANDA), B)
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=2, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    action = torch.rand(B, 3, device="cuda") * 2.4 - 1.2
    ctrl.set_action(action)
    ref = O.controller_step(fr, rows, action.cpu().numpy(), pos_only=True, **oracle_kw(ctrl))
    assert np.abs(art.drive_targets[:, :7].cpu().numpy() - ref["target_qpos"]).max() < Q_TOL


def test_fetch_masked_delta_pose_with_uncontrolled_ancestors():
    B = 512
    art, ctrl, fr = controller(FETCH, pose_cfg(FETCH), B)
    k = ctrl.kinematics
    assert k.chain_table.n == 11 and k.n_ctrl == 7 and k.kernel_name == "fetch_gripper"
    q = sample_q(fr, B, seed=3, shrink=0.3)
    qa, rows = fill_qpos(art, fr, k.active_ancestor_joint_idxs, q)
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    tgt = art.px.cuda_articulation_target_qpos.torch()
    tgt.fill_(9.0)
    action = torch.rand(B, 6, device="cuda") * 2 - 1
    ctrl.set_action(action)
    ref = O.controller_step(fr, rows, action.cpu().numpy(), **oracle_kw(ctrl))
    cols = ctrl.active_joint_indices.long()
    got = art.drive_targets[:, cols].cpu().numpy()
    mask = np.zeros(11, bool)
    mask[k.controlled_joints_idx_in_qmask] = True
    ok = well(fr, qa, mask)
    assert ok.sum() > 50 and np.abs(got - ref["target_qpos"])[ok].max() < Q_TOL
    others = [c for c in range(art.max_dof) if c not in cols.tolist()]
    assert (art.drive_targets[:, others] == 9.0).all()  # base / torso / head / gripper targets untouched


@pytest.mark.parametrize("kind", ["pose", "pos"])
def test_use_target_delta_iterative(kind):
    """pd_ee_target_delta_pose / _pos: target pose is controller state, iterative solve."""
    B = 600
    cfg = pose_cfg(PANDA, use_target=True) if kind == "pose" else pos_cfg(PANDA, use_target=True)
    art, ctrl, fr = controller(PANDA, cfg, B)
    nd = 6 if kind == "pose" else 3
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=4, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    prev = ctrl._target_pose.raw_pose.cpu().numpy().astype(np.float64)
    fkp = ctrl.kinematics.forward_kinematics(art.get_qpos()).raw_pose.cpu().numpy()
    assert np.abs(prev[:, :3] - fkp[:, :3]).max() < 1e-6  # reset: target pose = ee pose at base
    for step in range(2):
        action = torch.rand(B, nd, device="cuda") * 2 - 1
        ctrl.set_action(action)
        ref = O.controller_step(fr, rows, action.cpu().numpy(), use_target=True, pos_only=(kind == "pos"), prev_target_pose=prev,
                                **oracle_kw(ctrl))
        tp = ctrl.get_state()["target_pose"].cpu().numpy()
        assert np.abs(tp[:, :3] - ref["target_pose"][:, :3]).max() < 1e-6
        err = np.abs(art.drive_targets[:, :7].cpu().numpy() - ref["target_qpos"]).max(1)
        assert (err > Q_TOL).mean() < 0.02 and np.median(err) < 2e-6
        prev = ref["target_pose"]  # target accumulates; qpos unchanged (no sim step in the shim)
    st = ctrl.get_state()
    ctrl.set_state({"target_pose": st["target_pose"].clone()})
    assert torch.equal(ctrl.get_state()["target_pose"], st["target_pose"])


def test_absolute_pd_ee_pose():
    B = 400
    cfg = pose_cfg(PANDA, pos_lower=None, pos_upper=None, rot_lower=None, rot_upper=None, use_delta=False, normalize_action=False)
    art, ctrl, fr = controller(PANDA, cfg, B)
    q = sample_q(fr, B, seed=5, rest=PANDA_REST)
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, q)
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    # reachable absolute targets: FK of a perturbed state, expressed as xyz + euler XYZ
    q2 = qa + np.random.default_rng(6).uniform(-0.1, 0.1, qa.shape)
    _, R2, p2 = O.jacobian(fr, q2)
    sy = np.clip(R2[:, 0, 2], -1, 1)
    e1 = np.arcsin(sy)
    e0 = np.arctan2(-R2[:, 1, 2], R2[:, 2, 2])
    e2 = np.arctan2(-R2[:, 0, 1], R2[:, 0, 0])
    euler = np.stack([e0, e1, e2], 1)
    assert np.allclose(O.euler_xyz_to_matrix(euler), R2, atol=1e-9)
    action = torch.from_numpy(np.concatenate([p2, euler], 1).astype(np.float32)).cuda()
    ctrl.set_action(action)
    ref = O.controller_step(fr, rows, action.cpu().numpy().astype(np.float64), use_delta=False, normalize_action=False, **oracle_kw(ctrl))
    err = np.abs(art.drive_targets[:, :7].cpu().numpy() - ref["target_qpos"]).max(1)
    assert (err > Q_TOL).mean() < 0.02 and np.median(err) < 2e-6
    full = qa.copy()
    full[:, :7] = art.drive_targets[:, :7].cpu().numpy()
    _, _, ps = O.jacobian(fr, full)
    assert np.linalg.norm(ps - p2, axis=1).max() < 1.2e-3


def test_interpolate_and_before_simulation_step():
    B = 300
    art, ctrl, fr = controller(PANDA, pose_cfg(PANDA, interpolate=True), B)
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=7, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    tgt = art.px.cuda_articulation_target_qpos.torch()
    tgt.fill_(5.0)
    action = torch.rand(B, 6, device="cuda") * 2 - 1
    ctrl.set_action(action)
    assert (tgt == 5.0).all()  # interpolate: set_action does not write drive targets (pd_ee_pose.py:112-115)
    ref = O.controller_step(fr, rows, action.cpu().numpy(), **oracle_kw(ctrl))["target_qpos"]
    start = rows[:, :7].astype(np.float64)
    for k in range(1, 6):
        ctrl.before_simulation_step()
        want = start + (ref - start) / 5 * k
        assert np.abs(art.drive_targets[:, :7].cpu().numpy() - want).max() < Q_TOL
    assert (art.drive_targets[:, 7:] == 5.0).all()


def test_reset_is_masked():
    B = 64
    art, ctrl, fr = controller(PANDA, pose_cfg(PANDA, use_target=True), B)
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=8, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    ctrl.reset()
    tp0 = ctrl._target_pose.raw_pose.clone()
    sq0 = ctrl._start_qpos.clone()
    # move every env, reset only the even ones
    qa2, rows2 = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=9, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    art.scene._reset_mask[:] = False
    art.scene._reset_mask[::2] = True
    ctrl.reset()
    tp1, sq1 = ctrl._target_pose.raw_pose, ctrl._start_qpos
    assert torch.equal(tp1[1::2], tp0[1::2]) and torch.equal(sq1[1::2], sq0[1::2])
    new_pose = ctrl.ee_pose_at_base.raw_pose
    assert torch.equal(tp1[::2], new_pose[::2])
    assert np.abs(sq1[::2].cpu().numpy() - rows2[::2, :7]).max() == 0
    assert torch.equal(ctrl._target_qpos[::2], sq1[::2])


def test_partial_reset_after_lazy_set_action_keeps_running_envs():
    """in lazy state mode `set_action` leaves the dense qpos copies un-materialised; the next
    PARTIAL reset must still update only the masked envs -- in particular the accumulated `_target_pose` of the
    envs that keep running (use_target, pd_ee_pose.py:66-74)."""
    B = 96
    art, ctrl, fr = controller(PANDA, pose_cfg(PANDA, use_target=True), B)
    assert ctrl.state_mode == "lazy"
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=21, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    art.scene._reset_mask[:] = True
    ctrl.reset()
    ctrl.set_action(torch.rand(B, 6, device="cuda") * 2 - 1)
    tp_run = ctrl._target_pose.raw_pose.clone()      # accumulated target pose after one step
    tq_run = art.drive_targets[:, :7].clone()        # what the step wrote
    assert not torch.equal(tp_run, ctrl.ee_pose_at_base.raw_pose)
    # the simulator moves every env; only every third env is reset
    qa2, rows2 = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=22, rest=PANDA_REST))
    set_link_poses_from_fk(art, ctrl)
    art.scene._reset_mask[:] = False
    art.scene._reset_mask[::3] = True
    ctrl.reset()
    keep = ~art.scene._reset_mask
    ee = ctrl.ee_pose_at_base.raw_pose
    tp = ctrl._target_pose.raw_pose
    assert torch.equal(tp[keep], tp_run[keep]), "a partial reset overwrote the target pose of running envs"
    assert torch.equal(tp[~keep], ee[~keep])
    assert torch.equal(ctrl._target_qpos[keep], tq_run[keep])
    assert np.abs(ctrl._target_qpos[~keep].cpu().numpy() - rows2[::3, :7]).max() == 0
    assert np.abs(ctrl._start_qpos[~keep].cpu().numpy() - rows2[::3, :7]).max() == 0


def test_body_rows_are_int64_like_the_reference_and_converted_once():
    """`_body_data_index` is int64 in the reference (structs/base.py:104-110); the int32 copies
    the launch reads must outlive the call (no pointer to a temporary) and must not alias each other."""
    B = 300
    art, ctrl, fr = controller(PANDA, pose_cfg(PANDA), B)
    assert art.root._body_data_index.dtype == torch.int64 and ctrl.ee_link._body_data_index.dtype == torch.int64
    rng = np.random.default_rng(5)
    quat = rng.normal(size=(B, 4))
    quat /= np.linalg.norm(quat, axis=1, keepdims=True)
    root = np.concatenate([rng.uniform(-1, 1, (B, 3)), quat], 1)
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=12))
    set_link_poses_from_fk(art, ctrl, root)
    r0, e0 = ctrl._body_rows()
    got = ctrl.ee_pose_at_base.raw_pose
    r1, e1 = ctrl._body_rows()
    assert r0.data_ptr() == r1.data_ptr() and e0.data_ptr() == e1.data_ptr() and r0.data_ptr() != e0.data_ptr()
    want = (art.pose.inv() * ctrl.ee_pose).raw_pose
    qd = torch.minimum((got[:, 3:] - want[:, 3:]).abs().amax(1), (got[:, 3:] + want[:, 3:]).abs().amax(1))
    assert (got[:, :3] - want[:, :3]).abs().max() < 2e-6 and qd.max() < 2e-6


def test_ee_pose_at_base_kernel_matches_pose_algebra():
    B = 500
    art, ctrl, fr = controller(PANDA, pose_cfg(PANDA), B)
    rng = np.random.default_rng(10)
    quat = rng.normal(size=(B, 4))
    quat /= np.linalg.norm(quat, axis=1, keepdims=True)
    root = np.concatenate([rng.uniform(-1, 1, (B, 3)), quat], 1)
    qa, rows = fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=11))
    set_link_poses_from_fk(art, ctrl, root)
    got = ctrl.ee_pose_at_base.raw_pose.cpu().numpy()
    ee_world = art.links_map[PANDA[1]].pose.raw_pose.cpu().numpy().astype(np.float64)
    want = O.ee_pose_at_base(root.astype(np.float32).astype(np.float64), ee_world)
    assert np.abs(got[:, :3] - want[:, :3]).max() < 2e-6
    assert np.minimum(np.abs(got[:, 3:] - want[:, 3:]).max(1), np.abs(got[:, 3:] + want[:, 3:]).max(1)).max() < 2e-6
    # and it equals FK in the root frame (what PhysX would report)
    _, R, p = O.jacobian(fr, qa)
    assert np.abs(got[:, :3] - p).max() < 5e-6


def test_combined_controller_arm_plus_gripper_flat_action():
    B = 256
    art = make_art(PANDA, B)
    arm = pose_cfg(PANDA)
    grip = PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01, upper=0.04,
                                           stiffness=1e3, damping=1e2, force_limit=100)
    cc = CombinedController(dict(arm=arm, gripper=grip), art, control_freq=20, sim_freq=100, scene=art.scene)
    assert cc.action_space.shape == (B, 7) and cc.action_mapping == {"arm": (0, 6), "gripper": (6, 7)}
    cc.set_drive_property()
    assert art.joints_map["panda_joint3"].drive_properties == (1e3, 1e2, 100.0, "force")
    fr = O.load_chain(PANDA[0], PANDA[1])
    armc = cc.controllers["arm"]
    qa, rows = fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=12, rest=PANDA_REST))
    set_link_poses_from_fk(art, armc)
    cc.reset()
    action = torch.rand(B, 7, device="cuda") * 2 - 1
    cc.set_action(action)
    ref = O.controller_step(fr, rows, action[:, :6].cpu().numpy(), **oracle_kw(armc))
    got = art.drive_targets.cpu().numpy()
    assert np.abs(got[:, :7] - ref["target_qpos"]).max() < Q_TOL
    g = 0.5 * (0.04 - 0.01) + 0.5 * (0.04 + 0.01) * np.clip(action[:, 6].cpu().numpy(), -1, 1)
    assert np.abs(got[:, 7] - g).max() < 1e-6 and np.abs(got[:, 8] - g).max() < 1e-6
    for _ in range(5):
        cc.before_simulation_step()
    with pytest.raises(AssertionError):
        cc.set_action(action[:, :6])


def _combined(robot, B, extra, fuse, **artkw):
    from maniskill_b200 import _lib as L

    art = make_art(robot, B, **artkw)
    cfgs = dict(arm=pose_cfg(robot))
    cfgs.update(extra)
    cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
    cc.fuse_siblings = fuse
    return art, cc, L.load()


PANDA_GRIP = dict(gripper=PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01,
                                                          upper=0.04, stiffness=1e3, damping=1e2, force_limit=100))


@pytest.mark.parametrize("B,artkw", [(4096, {}), (1000, {}), (333, dict(max_dof=11)), (257, dict(n_pad_rows=3))])
def test_fused_siblings_match_separate_controllers(B, artkw):
    """CombinedController with the gripper's columns fused into the arm's launch
    (msik_delta_step_fused) writes bit-identical drive targets to arm launch + eager gripper
    controller, in ONE launch; foreign columns / rows are preserved."""
    fr = O.load_chain(PANDA[0], PANDA[1])
    action = torch.rand(B, 7, device="cuda") * 3 - 1.5  # exercises the clip
    outs, launches = [], []
    for fuse in (False, True):
        art, cc, lib = _combined(PANDA, B, PANDA_GRIP, fuse, **artkw)
        armc = cc.controllers["arm"]
        fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=21, rest=PANDA_REST), seed=3)
        tgt = art.px.cuda_articulation_target_qpos.torch()
        tgt.copy_(torch.arange(tgt.numel(), device="cuda", dtype=torch.float32).reshape(tgt.shape))  # sentinel values
        set_link_poses_from_fk(art, armc)
        cc.reset()
        n0 = lib.msik_launch_count()
        cc.set_action(action)
        torch.cuda.synchronize()
        launches.append(lib.msik_launch_count() - n0)
        outs.append(tgt.clone())
        if fuse:
            g = cc.controllers["gripper"]
            assert torch.equal(g._target_qpos, art.drive_targets[:, 7:9])
            assert torch.equal(g._start_qpos, art.qpos[:, 7:9])
    # identity rows: everything in ONE launch; interleaved rows (gather variant): arm + one small sibling launch
    assert launches[1] == (2 if artkw.get("n_pad_rows") else 1), launches
    assert torch.equal(outs[0], outs[1])
    if "max_dof" in artkw:  # columns 9, 10 belong to nobody: sentinels survive
        ref = torch.arange(outs[1].numel(), device="cuda", dtype=torch.float32).reshape(outs[1].shape)
        assert torch.equal(outs[1][:, 9:], ref[:, 9:])


def test_fused_fetch_body_delta_and_gripper():
    """Fetch: arm (7 of 11 ancestors) + gripper mimic + body pd_joint_delta_pos (use_delta) fused;
    the base columns stay untouched (pd_base_vel writes velocity targets, fetch.py:100-113)."""
    from maniskill_b200.controllers import PDJointPosControllerConfig

    B = 2048
    extra = dict(
        gripper=PDJointPosMimicControllerConfig(joint_names=["l_gripper_finger_joint", "r_gripper_finger_joint"], lower=-0.01, upper=0.05,
                                                stiffness=1e3, damping=1e2, force_limit=100),
        body=PDJointPosControllerConfig(joint_names=["head_pan_joint", "head_tilt_joint", "torso_lift_joint"], lower=-0.1, upper=0.1,
                                        stiffness=1e3, damping=1e2, force_limit=100, use_delta=True),
    )
    fr = O.load_chain(FETCH[0], FETCH[1])
    action = torch.rand(B, 6 + 1 + 3, device="cuda") * 2 - 1
    outs = []
    for fuse in (False, True):
        art, cc, lib = _combined(FETCH, B, extra, fuse)
        armc = cc.controllers["arm"]
        fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=22), seed=4)
        tgt = art.px.cuda_articulation_target_qpos.torch()
        tgt.fill_(-7.0)
        set_link_poses_from_fk(art, armc)
        cc.reset()
        n0 = lib.msik_launch_count()
        cc.set_action(action)
        torch.cuda.synchronize()
        nl = lib.msik_launch_count() - n0
        outs.append(tgt.clone())
    assert nl == 1
    assert torch.equal(outs[0], outs[1])
    assert (outs[1] == -7.0).any(dim=0).sum() == 3  # the three base columns


def test_small_then_large_shared_memory_configuration_of_the_same_kernel():
    """Regression: the dynamic-shared-memory attribute is per FUNCTION; a NO_DRIVE call (small
    carve-up) followed by a drive call (larger carve-up) of the same kernel must both launch."""
    B = 4096
    fr = O.load_chain(PANDA[0], PANDA[1])
    art, cc, lib = _combined(PANDA, B, PANDA_GRIP, True)
    armc = cc.controllers["arm"]
    qa, rows = fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=31, rest=PANDA_REST))
    set_link_poses_from_fk(art, armc)
    cc.reset()
    a = torch.rand(B, 7, device="cuda") * 2 - 1
    h_out = armc.kinematics.delta_step_host(torch.from_numpy(rows).pin_memory(), a.cpu().pin_memory(),
                                            normalize=((-0.1,) * 3, (0.1,) * 3, -0.1))
    cc.set_action(a)
    torch.cuda.synchronize()
    assert torch.equal(h_out, art.drive_targets[:, :7].cpu())


def test_set_action_is_cuda_graph_capturable_and_stream_ordered():
    """The controller step launches on torch's current stream with no host sync, so a training
    loop can capture it in a CUDA graph; replay gives bit-identical drive targets."""
    B = 8192
    fr = O.load_chain(PANDA[0], PANDA[1])
    art, cc, lib = _combined(PANDA, B, PANDA_GRIP, True)
    armc = cc.controllers["arm"]
    fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=41, rest=PANDA_REST))
    set_link_poses_from_fk(art, armc)
    cc.reset()
    action = torch.rand(B, 7, device="cuda") * 2 - 1
    cc.set_action(action)  # warm-up outside capture (first-launch attribute / occupancy queries)
    torch.cuda.synchronize()
    want = art.px.cuda_articulation_target_qpos.torch().clone()
    art.px.cuda_articulation_target_qpos.torch().zero_()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        cc.set_action(action)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        cc.set_action(action)
    art.px.cuda_articulation_target_qpos.torch().zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(art.px.cuda_articulation_target_qpos.torch(), want)
    # new inputs in the same buffers are picked up by the replay
    action.mul_(0.5)
    g.replay()
    torch.cuda.synchronize()
    got = art.px.cuda_articulation_target_qpos.torch().clone()
    cc.set_action(action)
    torch.cuda.synchronize()
    assert torch.equal(art.px.cuda_articulation_target_qpos.torch(), got)


@pytest.mark.parametrize("B,artkw", [(1, {}), (37, {}), (1000, {}), (4129, {}), (1000, dict(max_dof=11)), (515, dict(n_pad_rows=3))])
@pytest.mark.parametrize("fuse", [True, False])
def test_no_write_outside_the_target_rows(B, artkw, fuse):
    """compute-sanitizer is closed on this pool: own bounds check.  The drive-target buffer sits
    between two guard regions filled with a sentinel; after every kernel variant (TMA fused / merge,
    tiled, gather, ragged tails) the guards and all foreign rows / columns are untouched."""
    from maniskill_b200.shim import CudaArray

    fr = O.load_chain(PANDA[0], PANDA[1])
    art, cc, lib = _combined(PANDA, B, PANDA_GRIP, fuse, **artkw)
    n_rows, stride = art.px.cuda_articulation_target_qpos.torch().shape
    guard = 4096  # floats, keeps the 16-byte alignment of the tile bases
    big = torch.full((2 * guard + n_rows * stride,), -12345.0, device="cuda")
    art.px.cuda_articulation_target_qpos = CudaArray(big[guard:guard + n_rows * stride].view(n_rows, stride))
    armc = cc.controllers["arm"]
    fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=51, rest=PANDA_REST))
    set_link_poses_from_fk(art, armc)
    cc.reset()
    a = torch.rand(B, 7, device="cuda") * 2 - 1
    cc.set_action(a)
    torch.cuda.synchronize()
    assert (big[:guard] == -12345.0).all() and (big[-guard:] == -12345.0).all()
    tgt = art.px.cuda_articulation_target_qpos.torch()
    own = torch.zeros(n_rows, dtype=torch.bool, device="cuda")
    own[art._data_index.long()] = True
    assert (tgt[~own] == -12345.0).all()  # rows of other articulations
    assert (tgt[own][:, 9:] == -12345.0).all()  # columns beyond the robot's dofs
    assert torch.isfinite(tgt[own][:, :9]).all() and (tgt[own][:, :9] != -12345.0).all()


def test_realman_dual_arm_two_ee_controllers_in_one_combined_controller():
    """The fork's mobile dual-arm robot (agents/robots/realman/mobile_realman.py:298-350): right and left
    pd_ee_delta_pose controllers share one articulation and one flat action; each drives the 6 arm joints of its
    own chain (3 uncontrolled base joints first) and leaves every other column alone."""
    from conftest import REALMAN_L, REALMAN_R
    from maniskill_b200.shim import FakeArticulation

    B = 1024
    art = FakeArticulation(REALMAN_R[0], B, ee_link_names=[REALMAN_R[1], REALMAN_L[1]])
    cfgs = dict(right_arm=pose_cfg(REALMAN_R), left_arm=pose_cfg(REALMAN_L))
    cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
    right, left = cc.controllers["right_arm"], cc.controllers["left_arm"]
    assert right.kinematics.kernel_name == "realman_r_link6" and left.kinematics.kernel_name == "realman_l_link6"
    assert right.kinematics.chain_table.n == 9 and right.kinematics.n_ctrl == 6
    fr_r, fr_l = O.load_chain(*REALMAN_R[:2]), O.load_chain(*REALMAN_L[:2])
    rng = np.random.default_rng(21)
    rows = rng.uniform(-1.5, 1.5, (B, art.max_dof)).astype(np.float32)
    rows[:, :2] = rng.uniform(-0.1, 0.1, (B, 2))  # planar base inside its limits
    buf = art.px.cuda_articulation_qpos.torch()
    buf.normal_()
    buf[art._data_index.long()] = torch.from_numpy(rows).cuda()
    root = Pose.create(torch.tensor([[0.0, 0, 0, 1, 0, 0, 0]], device="cuda").repeat(B, 1))
    art.set_root_pose(root)
    for c in (right, left):
        art.set_link_pose(c.ee_link.name, root * c.kinematics.forward_kinematics(art.get_qpos()))
    cc.reset()
    tgt = art.px.cuda_articulation_target_qpos.torch()
    tgt.fill_(9.0)
    action = torch.rand(B, 12, device="cuda") * 2 - 1
    cc.set_action(action)
    touched = []
    for c, fr, a in ((right, fr_r, action[:, :6]), (left, fr_l, action[:, 6:])):
        k = c.kinematics
        ref = O.controller_step(fr, rows, a.cpu().numpy(), **oracle_kw(c))
        cols = c.active_joint_indices.long()
        touched += cols.tolist()
        got = art.drive_targets[:, cols].cpu().numpy()
        mask = np.zeros(9, bool)
        mask[k.controlled_joints_idx_in_qmask] = True
        qa = rows[:, np.asarray(k.active_ancestor_joint_idxs.cpu() if hasattr(k.active_ancestor_joint_idxs, "cpu") else k.active_ancestor_joint_idxs)].astype(np.float64)
        ok = well(fr, qa, mask)
        assert ok.sum() > 100 and np.abs(got - ref["target_qpos"])[ok].max() < Q_TOL
    others = [c for c in range(art.max_dof) if c not in touched]
    assert len(touched) == 12 and (art.drive_targets[:, others] == 9.0).all()


def test_fast_host_path_and_capture_match_the_general_path():
    """CombinedController.set_action's prebuilt-argument fast path (PDEEPosController.fast_step) and the CUDA-graph
    replay of BaseController.capture() write bit-identical drive targets to the general path, for contiguous actions;
    strided / sliced actions fall back to the general path; state read after a replay is this step's."""
    B = 4096
    fr = O.load_chain(PANDA[0], PANDA[1])
    art, cc, lib = _combined(PANDA, B, PANDA_GRIP, True)
    arm = cc.controllers["arm"]
    fill_qpos(art, fr, arm.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=31, rest=PANDA_REST), seed=5)
    set_link_poses_from_fk(art, arm)
    cc.reset()
    tgt = art.px.cuda_articulation_target_qpos.torch()
    action = torch.rand(B, 7, device="cuda") * 2 - 1
    assert cc._program() is not None
    tgt.fill_(-7.0)
    cc._set_action_general(action)
    want = tgt.clone()
    tgt.fill_(-7.0)
    n0 = lib.msik_launch_count()
    cc.set_action(action)  # fast path
    assert lib.msik_launch_count() - n0 == 1
    assert torch.equal(tgt, want)
    # a column slice of a wider buffer is not the fast path's layout: same result through the fallback
    wide = torch.zeros(B, 9, device="cuda")
    wide[:, 1:8] = action
    tgt.fill_(-7.0)
    cc.set_action(wide[:, 1:8])
    assert torch.equal(tgt, want)
    # graph replay with the action refilled in place
    replay = cc.capture(action)
    action.copy_(torch.rand(B, 7, device="cuda") * 2 - 1)
    tgt.fill_(-7.0)
    cc._set_action_general(action)
    want2 = tgt.clone()
    tgt.fill_(-7.0)
    replay()
    torch.cuda.synchronize()
    assert torch.equal(tgt, want2)
    assert torch.equal(arm._target_qpos, want2[:, :7])
    assert torch.equal(cc.controllers["gripper"]._target_qpos, want2[:, 7:9])


def _realman_dict(kind):
    """Controller dicts of the fork's Realman robot (agents/robots/realman/mobile_realman.py:298-443)."""
    from conftest import REALMAN_L, REALMAN_R
    from maniskill_b200 import presets

    if kind == "stock_pos":  # the one EE mode the fork ships enabled: pd_ee_delta_pos (mobile_realman.py:436-443)
        return presets.realman_pd_ee_delta_pos(REALMAN_R[0])
    mk = pose_cfg if kind == "pose" else pos_cfg
    return dict(right_arm=mk(REALMAN_R), left_arm=mk(REALMAN_L))


@pytest.mark.parametrize("kind,B", [("pose", 1000), ("pose", 4099), ("pos", 2049), ("stock_pos", 4096), ("stock_pos", 777)])
def test_realman_both_arms_in_one_launch(kind, B):
    """msik_delta_step_dual: the second EE controller of the CombinedController is solved inside the lead's persistent
    launch (one pass over the 164-byte rows).  Same drive targets / velocity targets as the controllers run one by one
    (fuse_siblings = False), ONE launch, through the general path and the prepared-step fast path; eager state mode falls
    back to the second arm's own launch with the same result."""
    from conftest import REALMAN_L, REALMAN_R
    from maniskill_b200 import _lib as L
    from maniskill_b200.shim import FakeArticulation

    lib = L.load()
    rng = np.random.default_rng(5)
    outs = {}
    for mode in ("separate", "general", "fast", "eager"):
        art = FakeArticulation(REALMAN_R[0], B, ee_link_names=[REALMAN_R[1], REALMAN_L[1]])
        cc = CombinedController(_realman_dict(kind), art, control_freq=20, sim_freq=100, scene=art.scene)
        cc.fuse_siblings = mode != "separate"
        if mode == "eager":
            cc.controllers["left_arm"].state_mode = "eager"
        rows = np.random.default_rng(6).uniform(-1.5, 1.5, (B, art.max_dof)).astype(np.float32)
        rows[:, :2] = np.random.default_rng(7).uniform(-0.1, 0.1, (B, 2))
        art.px.cuda_articulation_qpos.torch().copy_(torch.from_numpy(rows).cuda())
        root = Pose.create(torch.tensor([[0.0, 0, 0, 1, 0, 0, 0]], device="cuda").repeat(B, 1))
        art.set_root_pose(root)
        for u in ("right_arm", "left_arm"):
            c = cc.controllers[u]
            art.set_link_pose(c.ee_link.name, root * c.kinematics.forward_kinematics(art.get_qpos()))
        cc.reset()
        tgt = art.px.cuda_articulation_target_qpos.torch()
        qvel = art.px.cuda_articulation_target_qvel.torch()
        tgt.fill_(9.0)
        qvel.fill_(-5.0)
        adim = cc.single_action_space.shape[0]
        action = torch.from_numpy(np.random.default_rng(8).uniform(-1.2, 1.2, (B, adim)).astype(np.float32)).cuda()
        n0 = lib.msik_launch_count()
        if mode == "general":
            cc._set_action_general(action)
        else:
            cc.set_action(action)
        torch.cuda.synchronize()
        nl = lib.msik_launch_count() - n0
        if mode in ("general", "fast"):
            assert nl == 1, (mode, nl)
            assert cc._fusion_plan()["dual"] is not None
        outs[mode] = (tgt.clone(), qvel.clone())
        if mode == "fast":  # lazily gathered state of BOTH arms is this step's
            for u in ("right_arm", "left_arm"):
                c = cc.controllers[u]
                assert torch.equal(c._target_qpos, tgt[:, c.active_joint_indices.long()])
    ref_t, ref_v = outs["separate"]
    touched = (ref_t != 9.0).any(dim=0)
    assert touched.sum().item() == (12 if kind != "stock_pos" else 12 + 12 + 2)
    for mode in ("general", "fast", "eager"):
        t, v = outs[mode]
        assert (t - ref_t).abs().max().item() <= 1e-6, mode
        assert torch.equal(t[:, ~touched], ref_t[:, ~touched]), mode
        assert torch.equal(v, ref_v), mode
    assert torch.equal(outs["general"][0], outs["fast"][0])


@pytest.mark.parametrize("robot", ["panda", "realman"])
def test_prepared_step_with_an_unaligned_action_buffer(robot):
    """The prepared step is planned for a 16-byte aligned action buffer; an action tensor that starts 4 bytes into an
    allocation (contiguous, so it takes the fast path) must re-plan inside msik_step_launch -- bulk copies need aligned
    rows, so the tiled kernels serve it (for Realman: the second arm's own launch) -- with the same drive targets."""
    from conftest import REALMAN_L, REALMAN_R
    from maniskill_b200 import presets
    from maniskill_b200.shim import FakeArticulation

    B = 1000
    if robot == "panda":
        art, cc, lib = _combined(PANDA, B, PANDA_GRIP, True)
        ees = [cc.controllers["arm"]]
    else:
        art = FakeArticulation(REALMAN_R[0], B, ee_link_names=[REALMAN_R[1], REALMAN_L[1]])
        cc = CombinedController(presets.realman_pd_ee_delta_pos(REALMAN_R[0]), art, control_freq=20, sim_freq=100, scene=art.scene)
        ees = [cc.controllers["right_arm"], cc.controllers["left_arm"]]
    g = torch.Generator(device="cuda").manual_seed(3)
    q = art.px.cuda_articulation_qpos.torch()
    q.copy_(torch.rand(q.shape, generator=g, device="cuda") * 2 - 1)
    root = Pose.create(torch.tensor([[0.0, 0, 0, 1, 0, 0, 0]], device="cuda").repeat(B, 1))
    art.set_root_pose(root)
    for c in ees:
        art.set_link_pose(c.ee_link.name, root * c.kinematics.forward_kinematics(art.get_qpos()))
    cc.reset()
    adim = cc.single_action_space.shape[0]
    aligned = torch.rand((B, adim), generator=g, device="cuda") * 2 - 1
    store = torch.empty(B * adim + 1, device="cuda")
    shifted = store[1:].view(B, adim)
    shifted.copy_(aligned)
    assert shifted.is_contiguous() and shifted.data_ptr() % 16 == 4
    tgt = art.px.cuda_articulation_target_qpos.torch()
    qvel = art.px.cuda_articulation_target_qvel.torch()
    tgt.fill_(3.0); qvel.fill_(-2.0)
    cc.set_action(aligned)
    want_t, want_v = tgt.clone(), qvel.clone()
    tgt.fill_(3.0); qvel.fill_(-2.0)
    cc.set_action(shifted)
    torch.cuda.synchronize()
    assert (tgt - want_t).abs().max().item() <= 1e-6
    assert torch.equal(qvel, want_v)
    assert torch.equal(tgt == 3.0, want_t == 3.0)


def test_dependent_launch_reads_the_rows_the_previous_step_wrote():
    """The persistent kernel is scheduled while the previous kernel of the stream drains and prefetches its first rows into L2
    BEFORE griddepcontrol.wait (msik_tma.cuh).  A step that re-reads target rows (an arm-only controller: foreign columns are
    merged from the old rows) right after a step that rewrote those rows (the stock dict: whole rows) must see the new values.
    Eight such pairs inside ONE CUDA graph at 2^18 envs, with different gripper actions per pair: after every pair the gripper
    columns are the pair's own and the arm columns are the arm-only controller's."""
    B = 1 << 18
    fr = O.load_chain(PANDA[0], PANDA[1])
    art, cc, lib = _combined(PANDA, B, PANDA_GRIP, True)
    armc = cc.controllers["arm"]
    fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, 4096, seed=43, rest=PANDA_REST)[np.arange(B) % 4096])
    set_link_poses_from_fk(art, armc)
    cc.reset()
    solo_cfg = pose_cfg(PANDA)
    solo = solo_cfg.controller_cls(solo_cfg, art, control_freq=20, sim_freq=100, scene=art.scene)  # same buffers, arm columns only
    solo.reset()
    g = torch.Generator(device="cuda").manual_seed(17)
    n = 8
    acts = [torch.rand((B, 7), generator=g, device="cuda") * 2 - 1 for _ in range(n)]
    b = torch.rand((B, 6), generator=g, device="cuda") * 2 - 1
    tgt = art.px.cuda_articulation_target_qpos.torch()
    # expected pieces, each step alone
    grip = []
    for a in acts:
        cc.set_action(a)
        torch.cuda.synchronize()
        grip.append(tgt[:, 7:9].clone())
    assert not torch.equal(grip[0], grip[1])
    solo.set_action(b)
    torch.cuda.synchronize()
    arm_cols = tgt[:, :7].clone()
    snaps = [torch.empty_like(tgt) for _ in range(n)]
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        with torch.cuda.graph(gr, stream=s):
            for i in range(n):
                cc.set_action(acts[i])   # whole rows: arm + gripper columns
                solo.set_action(b)       # re-reads the rows just written, replaces the arm columns
                snaps[i].copy_(tgt)
    for rep in range(3):
        tgt.fill_(7.0)
        gr.replay()
        torch.cuda.synchronize()
        for i in range(n):
            assert torch.equal(snaps[i][:, 7:9], grip[i]), (rep, i)
            assert torch.equal(snaps[i][:, :7], arm_cols), (rep, i)


def test_iterative_regrouping_with_tiled_and_gathered_rows():
    """The regrouping instantiation of the iterative kernel (large batches) writes its results by each env's ROW, whatever lane
    finished it: absolute-pose steps at 40 000 envs through the CTA-tile path (identity row index) and through the per-lane
    gather path (an articulation whose rows are interleaved with padding rows) give the same drive targets as 5 000-env
    batches of the same states (no regrouping there), bit for bit."""
    cfg = pose_cfg(PANDA, pos_lower=None, pos_upper=None, rot_lower=None, rot_upper=None, use_delta=False, normalize_action=False)
    B, small = 40000, 5000
    fr = O.load_chain(PANDA[0], PANDA[1])
    q = sample_q(fr, small, seed=31, rest=PANDA_REST)
    q2 = q[:, :7] + np.random.default_rng(32).uniform(-1, 1, (small, 1)) * np.random.default_rng(33).uniform(-0.2, 0.2, (small, 7))
    qq = q.copy(); qq[:, :7] = q2
    _, R2, p2 = O.jacobian(fr, qq)
    e1 = np.arcsin(np.clip(R2[:, 0, 2], -1, 1))
    e0 = np.arctan2(-R2[:, 1, 2], R2[:, 2, 2])
    e2 = np.arctan2(-R2[:, 0, 1], R2[:, 0, 0])
    act_small = np.concatenate([p2, np.stack([e0, e1, e2], 1)], 1).astype(np.float32)
    want = None
    for artkw in ({}, dict(n_pad_rows=3)):
        outs = {}
        for n in (small, B):
            art, ctrl, _ = controller(PANDA, cfg, n, **artkw)
            rep = np.arange(n) % small
            fill_qpos(art, fr, ctrl.kinematics.active_ancestor_joint_idxs, q[rep])
            set_link_poses_from_fk(art, ctrl)
            ctrl.reset()
            ctrl.set_action(torch.from_numpy(act_small[rep]).cuda())
            torch.cuda.synchronize()
            outs[n] = art.drive_targets[:, :7].clone()
        assert torch.equal(outs[B][:small], outs[small]), artkw
        assert torch.equal(outs[B], outs[small][torch.arange(B, device="cuda") % small]), artkw
        if want is None:
            want = outs[B]
        else:
            assert torch.equal(outs[B], want)  # gathered rows == tiled rows
