"""pd_ee_delta_pos (3-dof action) step time at 2^20 Panda envs: CombinedController(arm pd_ee_delta_pos + mimic gripper).set_action (dev tool)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import PANDA
from helpers import PANDA_REST, make_art, pos_cfg, pose_cfg
from maniskill_b200.controllers import CombinedController, PDJointPosMimicControllerConfig

B = 1 << 20
for name, cfgf, adim in (("pd_ee_delta_pos", pos_cfg, 3), ("pd_ee_delta_pose", pose_cfg, 6)):
    groups = []
    for g in range(4):  # ring of 4 groups > L2
        art = make_art(PANDA, B)
        cfgs = dict(arm=cfgf(PANDA), gripper=PDJointPosMimicControllerConfig(joint_names=["panda_finger_joint1", "panda_finger_joint2"], lower=-0.01,
                                                                             upper=0.04, stiffness=1e3, damping=1e2, force_limit=100))
        cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
        rest = torch.tensor(PANDA_REST + [0.04, 0.04], dtype=torch.float32, device="cuda")
        art.px.cuda_articulation_qpos.torch().copy_(rest[None] + 0.02 * torch.randn((B, 9), device="cuda"))
        cc.reset()
        groups.append((cc, torch.rand((B, adim + 1), device="cuda") * 2 - 1))
    for cc, a in groups: cc.set_action(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    K = 200
    for i in range(K):
        cc, a = groups[i % 4]; cc.set_action(a)
    e1.record(); torch.cuda.synchronize()
    print(f"{name:18s} {e0.elapsed_time(e1) / K * 1e3:8.2f} us per 2^20 envs (eager API issue)")
    del groups
