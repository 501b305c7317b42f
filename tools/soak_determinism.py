"""Soak test for pipeline races (dev tool): repeat the same launches many times with L2-hot and L2-cold inputs and compare
every result bitwise with the first one.  Found nothing after the cross-proxy fence went in; found 10-100 wrong rows per
2^20 envs per launch before."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import FETCH, PANDA
from helpers import PANDA_REST, make_art, pose_cfg, pos_cfg
from maniskill_b200.controllers import CombinedController, PDJointPosMimicControllerConfig
from maniskill_b200.kinematics import Kinematics

REPS = int(os.environ.get("SOAK_REPS", "150"))
bad = 0
for B in ((1 << 20) + 37, 300000, 94720, 4096):
    # (1) compute_ik: NO_DRIVE + per-lane result stores, inputs L2-hot (same tensors every call)
    art = make_art(PANDA, B)
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in PANDA[2]], dtype=torch.int32, device="cuda")
    kin = Kinematics(PANDA[0], PANDA[1], art, idx)
    rest = torch.tensor(PANDA_REST + [0.04, 0.04], dtype=torch.float32, device="cuda")
    q = rest[None] + 0.02 * torch.randn((B, 9), device="cuda")
    a = (torch.rand((B, 6), device="cuda") * 2 - 1) * 0.1
    ref = kin.compute_ik(None, q, action=a, use_delta_ik_solver=True).clone()
    flush = torch.empty(64 << 20, device="cuda")
    for r in range(REPS):
        if r % 3 == 2: flush.fill_(float(r))  # every third call with a cold L2
        out = kin.compute_ik(None, q, action=a, use_delta_ik_solver=True)
        if not torch.equal(out, ref):
            bad += 1; print("compute_ik mismatch", B, r, (out != ref).any(dim=1).sum().item(), flush=True)
    # (2) the fused controller step (LEAN kernel, bulk stores of whole target rows) for 6-dof and 3-dof actions
    for cfgf, adim in ((pose_cfg, 6), (pos_cfg, 3)):
        cfgs = dict(arm=cfgf(PANDA), gripper=PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01,
                                                                             upper=0.04, stiffness=1e3, damping=1e2, force_limit=100))
        cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
        art.px.cuda_articulation_qpos.torch().copy_(q)
        cc.reset()
        act_t = torch.rand((B, adim + 1), device="cuda") * 2 - 1
        tgt = art.px.cuda_articulation_target_qpos.torch()
        cc.set_action(act_t); ref2 = tgt.clone()
        for r in range(REPS):
            if r % 3 == 2: flush.fill_(float(r))
            tgt.fill_(-3.0)
            cc.set_action(act_t)
            if not torch.equal(tgt, ref2):
                bad += 1; print("set_action mismatch", B, adim, r, (tgt != ref2).any(dim=1).sum().item(), flush=True)
    print("size", B, "done", flush=True)
    del art, kin, q, a, ref, flush
print("soak:", "CLEAN" if bad == 0 else f"{bad} mismatching launches", f"({REPS} repetitions per case)")
