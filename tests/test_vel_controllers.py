"""Velocity controllers that share a CombinedController with the EE controller:
pd_joint_vel / pd_base_vel / pd_base_forward_vel.  Golden vectors come from running the reference's own
set_action code (tools/gen_golden_vel.py); the oracle is held to them on the CPU, the CUDA kernel
(msik_vel_targets through the mirrored controller classes) on the GPU."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from conftest import FETCH
from oracle import oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "vel_golden.npz"))
BASE = ["root_x_axis_joint", "root_y_axis_joint", "root_z_rotation_joint"]
BASE_LOW, BASE_HIGH = [-1, -1, -3.14], [1, 1, 3.14]


def test_oracle_matches_reference_golden():
    ori = G["qpos"][:, 2].astype(np.float64)
    assert np.abs(O.base_vel_targets(ori, G["base_action"], BASE_LOW, BASE_HIGH) - G["base_target_qvel"]).max() < 1e-6
    assert np.abs(O.base_vel_targets(ori, G["base_action"], BASE_LOW, BASE_HIGH, normalize=False) - G["base_raw_target_qvel"]).max() < 1e-6
    assert np.abs(O.base_vel_targets(ori, G["fwd_action"], [-1, -3.14], [1, 3.14], forward_only=True) - G["fwd_target_qvel"]).max() < 1e-6
    assert np.abs(O.joint_vel_targets(G["joint_action"], [-1.0] * 7, [2.0] * 7) - G["joint_target_qvel"]).max() < 1e-6


def _art(B, **kw):
    from maniskill_b200.shim import FakeArticulation

    return FakeArticulation(FETCH[0], B, ee_link_names=[FETCH[1]], **kw)


def _fill(art, qpos_rows):
    buf = art.px.cuda_articulation_qpos.torch()
    buf.normal_()
    buf[art._data_index.long(), : qpos_rows.shape[1]] = torch.from_numpy(qpos_rows).cuda()
    tv = art.px.cuda_articulation_target_qvel.torch()
    tv.copy_(torch.arange(tv.numel(), device="cuda", dtype=torch.float32).reshape(tv.shape))  # sentinels
    return tv.clone()


@pytest.mark.gpu
@pytest.mark.parametrize("artkw", [{}, dict(n_pad_rows=3), dict(interleave_rows=True), dict(max_dof=17)])
def test_base_vel_matches_reference_golden(artkw):
    from maniskill_b200 import _lib as L
    from maniskill_b200.controllers import PDBaseVelControllerConfig

    B = G["qpos"].shape[0]
    art = _art(B, **artkw)
    for normalize, key in ((True, "base_target_qvel"), (False, "base_raw_target_qvel")):
        before = _fill(art, G["qpos"])
        cfg = PDBaseVelControllerConfig(joint_names=BASE, lower=BASE_LOW, upper=BASE_HIGH, damping=1000, force_limit=500,
                                        normalize_action=normalize)
        c = cfg.controller_cls(cfg, art, control_freq=20, sim_freq=100, scene=art.scene)
        n0 = L.load().msik_launch_count()
        c.set_action(torch.from_numpy(G["base_action"]).cuda())
        torch.cuda.synchronize()
        assert L.load().msik_launch_count() - n0 == 1
        after = art.px.cuda_articulation_target_qvel.torch()
        got = art.drive_velocity_targets[:, :3].cpu().numpy()
        assert np.abs(got - G[key]).max() < 1e-6
        # everything else in the PhysX buffer is untouched, bit for bit
        mask = torch.zeros_like(after, dtype=torch.bool)
        mask[art._data_index.long(), :3] = True
        assert torch.equal(after[~mask], before[~mask])


@pytest.mark.gpu
def test_forward_vel_and_joint_vel_match_reference_golden():
    from maniskill_b200.controllers import PDBaseForwardVelControllerConfig, PDJointVelControllerConfig

    B = G["qpos"].shape[0]
    art = _art(B)
    _fill(art, G["qpos"])
    cfg = PDBaseForwardVelControllerConfig(joint_names=BASE, lower=[-1, -3.14], upper=[1, 3.14], damping=1000, force_limit=500)
    c = cfg.controller_cls(cfg, art, control_freq=20, sim_freq=100, scene=art.scene)
    assert c.single_action_space.shape == (2,)
    c.set_action(torch.from_numpy(G["fwd_action"]).cuda())
    assert np.abs(art.drive_velocity_targets[:, :3].cpu().numpy() - G["fwd_target_qvel"]).max() < 1e-6
    names = [j.name for j in art.active_joints]
    cols = [int(x) for x in G["joint_cols"]]
    cfgj = PDJointVelControllerConfig(joint_names=[names[k] for k in cols], lower=-1.0, upper=2.0, damping=100, force_limit=100)
    cj = cfgj.controller_cls(cfgj, art, control_freq=20, sim_freq=100, scene=art.scene)
    before = art.px.cuda_articulation_target_qvel.torch().clone()
    cj.set_action(torch.from_numpy(G["joint_action"]).cuda())
    after = art.px.cuda_articulation_target_qvel.torch()
    assert torch.equal(after[:, cols].cpu(), torch.from_numpy(G["joint_target_qvel"]))  # elementwise, same roundings: bit-exact
    others = [k for k in range(after.shape[1]) if k not in cols]
    assert torch.equal(after[:, others], before[:, others])


@pytest.mark.gpu
@pytest.mark.parametrize("artkw", [{}, dict(n_pad_rows=2)])
def test_fetch_full_controller_dict_is_one_launch(artkw):
    """The stock Fetch pd_ee_delta_pose dict (fetch.py:273-278: arm + gripper + body + base): the arm's launch also
    writes gripper and body position targets AND the base controller's velocity targets (msik_delta_step_fused_vel);
    bit-identical to the controllers run one by one.  Gathered rows fall back to small extra launches inside the call."""
    from helpers import fill_qpos, pose_cfg, sample_q, set_link_poses_from_fk
    from maniskill_b200 import _lib as L
    from maniskill_b200.controllers import (CombinedController, PDBaseVelControllerConfig, PDJointPosControllerConfig,
                                            PDJointPosMimicControllerConfig)

    for B in (1, 33, 4099):
        outs = []
        for fuse in (True, False):
            art = _art(B, **artkw)
            cfgs = dict(
                arm=pose_cfg(FETCH),
                gripper=PDJointPosMimicControllerConfig(joint_names=["l_gripper_finger_joint", "r_gripper_finger_joint"], lower=-0.01, upper=0.05,
                                                        stiffness=1e3, damping=1e2, force_limit=100),
                body=PDJointPosControllerConfig(joint_names=["head_pan_joint", "head_tilt_joint", "torso_lift_joint"], lower=-0.1, upper=0.1,
                                                stiffness=1e3, damping=1e2, force_limit=100, use_delta=True),
                base=PDBaseVelControllerConfig(joint_names=BASE, lower=BASE_LOW, upper=BASE_HIGH, damping=1000, force_limit=500),
            )
            cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
            cc.fuse_siblings = fuse
            assert cc.single_action_space.shape == (6 + 1 + 3 + 3,)
            fr = O.load_chain(FETCH[0], FETCH[1])
            armc = cc.controllers["arm"]
            _, rows = fill_qpos(art, fr, armc.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=5), seed=6)
            set_link_poses_from_fk(art, armc)
            art.px.cuda_articulation_target_qvel.torch().fill_(-5.0)
            art.px.cuda_articulation_target_qpos.torch().fill_(-7.0)
            cc.reset()
            action = torch.rand(B, 13, device="cuda", generator=torch.Generator(device="cuda").manual_seed(B)) * 3 - 1.5
            n0 = L.load().msik_launch_count()
            cc.set_action(action)
            torch.cuda.synchronize()
            nl = L.load().msik_launch_count() - n0
            if fuse:
                assert nl == (1 if not artkw else 3), nl  # gathered rows: arm (gather kernel) + sibling pass + velocity pass
            ori = rows[:, 2].astype(np.float64)
            ref = O.base_vel_targets(ori, action[:, 10:].cpu().numpy(), BASE_LOW, BASE_HIGH)
            assert np.abs(art.drive_velocity_targets[:, :3].cpu().numpy() - ref).max() < 2e-6
            assert (art.drive_velocity_targets[:, 3:] == -5.0).all()
            outs.append((art.px.cuda_articulation_target_qpos.torch().clone(), art.px.cuda_articulation_target_qvel.torch().clone()))
        assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.gpu
def test_vel_targets_rejects_bad_configurations():
    from maniskill_b200 import _lib as L

    lib = L.load()
    cfg = L.VelCfg()
    cfg.n_envs, cfg.n_cols, cfg.ego = 4, 2, L.VEL_BASE  # pd_base_vel.py:15: at least x, y, orientation
    cfg.qpos_stride = cfg.qvel_stride = cfg.action_stride = 15
    buf = torch.zeros(4, 15, device="cuda")
    assert lib.msik_vel_targets(C.byref(cfg), buf.data_ptr(), None, buf.data_ptr(), buf.data_ptr(), None) != 0
    cfg.n_cols = 3
    cfg.col[0], cfg.col[1], cfg.col[2] = 0, 1, 1  # a column written twice
    assert lib.msik_vel_targets(C.byref(cfg), buf.data_ptr(), None, buf.data_ptr(), buf.data_ptr(), None) != 0
    cfg.col[2] = 2
    cfg.ori_col = 2
    cfg.action_offset = 13  # slice leaves the action row
    assert lib.msik_vel_targets(C.byref(cfg), buf.data_ptr(), None, buf.data_ptr(), buf.data_ptr(), None) != 0
    cfg.action_offset = 0
    assert lib.msik_vel_targets(C.byref(cfg), buf.data_ptr(), None, None, buf.data_ptr(), None) != 0
    assert lib.msik_vel_targets(C.byref(cfg), buf.data_ptr(), None, buf.data_ptr(), buf.data_ptr(), None) == 0
    torch.cuda.synchronize()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(12))
def test_vel_targets_random_layouts(seed):
    """msik_vel_targets with random row strides, column placement, action offset / stride, row_index gathers, ego modes,
    normalisation and ragged env counts against the oracle; cells the call must not touch are compared bit for bit."""
    from maniskill_b200 import _lib as L

    rng = np.random.default_rng(seed)
    lib = L.load()
    B = int(rng.choice([1, 31, 33, 257, 4099]))
    ego = int(rng.choice([L.VEL_JOINT, L.VEL_BASE, L.VEL_FORWARD]))
    n_cols = int(rng.integers(1 if ego == L.VEL_JOINT else 3, 9))
    qs = n_cols + int(rng.integers(1, 8))
    vs = qs + int(rng.integers(0, 3))
    cols = np.sort(rng.choice(qs, n_cols, replace=False)) if rng.random() < 0.5 else rng.permutation(qs)[:n_cols]
    na = n_cols - 1 if ego == L.VEL_FORWARD else n_cols
    aoff = int(rng.integers(0, 4))
    astride = aoff + na + int(rng.integers(0, 3))
    normalize = rng.random() < 0.6
    gather = rng.random() < 0.4
    M = B + int(rng.integers(0, 5)) if gather else B
    rows = rng.permutation(M)[:B].astype(np.int32) if gather else None
    ridx = rows if gather else np.arange(B)
    low = rng.uniform(-2, -0.5, na)
    high = rng.uniform(0.5, 2, na)
    qpos = rng.uniform(-3, 3, (M, qs)).astype(np.float32)
    action = rng.uniform(-1.4, 1.4, (B, astride)).astype(np.float32)
    vbuf = rng.normal(7, 1, (M, vs)).astype(np.float32)
    cfg = L.VelCfg()
    cfg.n_envs, cfg.n_cols, cfg.ego = B, n_cols, ego
    for k, c in enumerate(cols):
        cfg.col[k] = int(c)
    cfg.ori_col = int(cols[2]) if ego != L.VEL_JOINT else 0
    cfg.flags = L.AUX_NORMALIZE if normalize else 0
    cfg.qpos_stride, cfg.qvel_stride, cfg.action_stride, cfg.action_offset = qs, vs, astride, aoff
    lo32, hi32 = low.astype(np.float32), high.astype(np.float32)
    for k in range(na):
        cfg.mid[k] = float(np.float32(0.5) * (hi32[k] + lo32[k]))
        cfg.half[k] = float(np.float32(0.5) * (hi32[k] - lo32[k]))
    d_q, d_a, d_v = torch.from_numpy(qpos).cuda(), torch.from_numpy(action).cuda(), torch.from_numpy(vbuf).cuda()
    d_r = torch.from_numpy(rows).cuda() if gather else None
    rc = lib.msik_vel_targets(C.byref(cfg), d_q.data_ptr(), L.ptr(d_r), d_a.data_ptr(), d_v.data_ptr(), None)
    assert rc == 0, lib.msik_last_error()
    torch.cuda.synchronize()
    a = action[:, aoff:aoff + na].astype(np.float64)
    if ego == L.VEL_JOINT:
        ref = O.joint_vel_targets(a, lo32.astype(np.float64), hi32.astype(np.float64), normalize)
    else:
        ori = qpos[ridx, int(cols[2])].astype(np.float64)
        ref = O.base_vel_targets(ori, a, lo32.astype(np.float64), hi32.astype(np.float64), normalize, forward_only=(ego == L.VEL_FORWARD))
    got = d_v.cpu().numpy()
    assert np.abs(got[np.ix_(ridx, cols)] - ref).max() < 2e-6
    mask = np.zeros_like(vbuf, dtype=bool)
    mask[np.ix_(ridx, cols)] = True
    assert np.array_equal(got[~mask], vbuf[~mask])
