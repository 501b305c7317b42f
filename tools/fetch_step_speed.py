"""Stock Fetch CombinedController.set_action (arm pd_ee_delta_pose + gripper + body + base pd_base_vel, fetch.py:273-278):
everything fused into the arm's launch vs the controllers one by one (dev tool)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import FETCH
from helpers import fill_qpos, make_art, pose_cfg, sample_q, set_link_poses_from_fk
from maniskill_b200 import _lib as L
from maniskill_b200.controllers import (CombinedController, PDBaseVelControllerConfig, PDJointPosControllerConfig,
                                        PDJointPosMimicControllerConfig)
from oracle import oracle as O

BASE = ["root_x_axis_joint", "root_y_axis_joint", "root_z_rotation_joint"]
fr = O.load_chain(FETCH[0], FETCH[1])
for B in (4096, 1 << 20):
    for fuse in (True, False):
        groups = []
        for g in range(4 if B > 100000 else 1):
            art = make_art(FETCH, B)
            cfgs = dict(
                arm=pose_cfg(FETCH),
                gripper=PDJointPosMimicControllerConfig(joint_names=["l_gripper_finger_joint", "r_gripper_finger_joint"], lower=-0.01, upper=0.05,
                                                        stiffness=1e3, damping=1e2, force_limit=100),
                body=PDJointPosControllerConfig(joint_names=["head_pan_joint", "head_tilt_joint", "torso_lift_joint"], lower=-0.1, upper=0.1,
                                                stiffness=1e3, damping=1e2, force_limit=100, use_delta=True),
                base=PDBaseVelControllerConfig(joint_names=BASE, lower=[-1, -1, -3.14], upper=[1, 1, 3.14], damping=1000, force_limit=500),
            )
            cc = CombinedController(cfgs, art, control_freq=20, sim_freq=100, scene=art.scene)
            cc.fuse_siblings = fuse
            q = sample_q(fr, 4096, seed=g)
            fill_qpos(art, fr, cc.controllers["arm"].kinematics.active_ancestor_joint_idxs, q[np.random.default_rng(g).integers(0, 4096, B)], seed=g)
            cc.reset()
            groups.append((cc, torch.rand((B, 13), device="cuda") * 2 - 1))
        for cc, a in groups: cc.set_action(a)
        n0 = L.load().msik_launch_count()
        groups[0][0].set_action(groups[0][1])
        nl = L.load().msik_launch_count() - n0
        # CUDA graph of K steps: device time without the Python issue cost
        K = 40
        gr = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for i in range(3): groups[i % len(groups)][0].set_action(groups[i % len(groups)][1])
            torch.cuda.synchronize()
            with torch.cuda.graph(gr, stream=s):
                for i in range(K): groups[i % len(groups)][0].set_action(groups[i % len(groups)][1])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        gr.replay(); torch.cuda.synchronize()
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        t_graph = e0.elapsed_time(e1) / K * 1e3
        import time
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for i in range(K): groups[i % len(groups)][0].set_action(groups[i % len(groups)][1])
        torch.cuda.synchronize(); t_eager = (time.perf_counter() - t0) / K * 1e6
        print(f"Fetch full dict  B={B:8d}  fused={fuse!s:5s}  launches/step={nl}  graph {t_graph:7.2f} us/step   eager wall {t_eager:7.2f} us/step", flush=True)
        del groups
