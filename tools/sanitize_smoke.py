"""Small workload for `compute-sanitizer --tool memcheck` (one tool per gpurun call): every kernel
variant once on ragged sizes -- TMA path (fused / merge / NO_DRIVE), tiled, gather, iterative, FK."""
import os, sys
import numpy as np, torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import FETCH, PANDA
from helpers import PANDA_REST, fill_qpos, make_art, pose_cfg, sample_q, set_link_poses_from_fk
from maniskill_b200.controllers import CombinedController, PDJointPosMimicControllerConfig
from maniskill_b200.structs.pose import Pose
from oracle import oracle as O

def run(robot, B, fuse, **artkw):
    art = make_art(robot, B, **artkw)
    cfgs = dict(arm=pose_cfg(robot))
    if robot is PANDA:
        cfgs["gripper"] = PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01,
                                                          upper=0.04, stiffness=1e3, damping=1e2, force_limit=100)
    cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
    cc.fuse_siblings = fuse
    arm = cc.controllers["arm"]
    fr = O.load_chain(robot[0], robot[1])
    rest = PANDA_REST if robot is PANDA else None
    fill_qpos(art, fr, arm.kinematics.active_ancestor_joint_idxs, sample_q(fr, B, seed=5, rest=rest))
    set_link_poses_from_fk(art, arm)
    cc.reset()
    a = torch.rand(B, cc.action_space.shape[1] if B > 1 else cc.action_space.shape[0], device="cuda") * 2 - 1
    cc.set_action(a)
    q = art.get_qpos()
    pose, J = arm.kinematics.forward_kinematics(q, return_jacobian=True)
    arm.kinematics.compute_ik(pose, q)  # iterative
    arm.kinematics.compute_ik(None, q, action=a[:, :6].contiguous() * 0.1, use_delta_ik_solver=True)
    torch.cuda.synchronize()
    assert torch.isfinite(art.drive_targets).all()

for B in (1, 37, 1000):
    run(PANDA, B, True); run(PANDA, B, False); run(PANDA, B, True, max_dof=11); run(PANDA, B, True, n_pad_rows=3)
    run(FETCH, B, False)
os.environ["MSIK_DELTA_KERNEL"] = "tiled"
print("sanitize smoke ok")
