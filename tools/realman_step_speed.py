"""The fork's Realman dict (both arms pd_ee_delta_pos / pose [+ grippers + body + base]): both arms in ONE pass of the persistent
kernel (msik_delta_step_dual) vs one launch per arm (MSIK dual off: CombinedController.fuse_siblings = False keeps everything
separate; `lead only` = siblings fused but the second arm on its own).  CUDA-graph device time per step (dev tool)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import REALMAN_L, REALMAN_R
from test_gpu_controller import _realman_dict
from maniskill_b200 import _lib as L
from maniskill_b200.controllers import CombinedController
from maniskill_b200.shim import FakeArticulation
from maniskill_b200.structs.pose import Pose

for kind in ("stock_pos", "pose"):
    for B in (4096, 1 << 20):
        for mode in ("dual", "second arm separate"):
            groups = []
            for g in range(3 if B > 100000 else 1):
                art = FakeArticulation(REALMAN_R[0], B, ee_link_names=[REALMAN_R[1], REALMAN_L[1]])
                cc = CombinedController(_realman_dict(kind), art, control_freq=20, sim_freq=100, scene=art.scene)
                if mode != "dual":
                    plan = cc._fusion_plan()
                    plan["siblings"].discard("left_arm"); plan["dual"] = None
                art.px.cuda_articulation_qpos.torch().uniform_(-1.5, 1.5)
                root = Pose.create(torch.tensor([[0.0, 0, 0, 1, 0, 0, 0]], device="cuda").repeat(B, 1))
                art.set_root_pose(root)
                for u in ("right_arm", "left_arm"):
                    c = cc.controllers[u]
                    art.set_link_pose(c.ee_link.name, root * c.kinematics.forward_kinematics(art.get_qpos()))
                cc.reset()
                groups.append((cc, torch.rand((B, cc.single_action_space.shape[0]), device="cuda") * 2 - 1))
            for cc, a in groups: cc.set_action(a)
            n0 = L.load().msik_launch_count(); groups[0][0].set_action(groups[0][1]); nl = L.load().msik_launch_count() - n0
            K = 30
            gr = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for i in range(3): groups[i % len(groups)][0].set_action(groups[i % len(groups)][1])
                torch.cuda.synchronize()
                with torch.cuda.graph(gr, stream=s):
                    for i in range(K): groups[i % len(groups)][0].set_action(groups[i % len(groups)][1])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            gr.replay(); torch.cuda.synchronize()
            e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
            print(f"Realman {kind:9s} B={B:8d}  {mode:20s} launches/step={nl}  {e0.elapsed_time(e1) / K * 1e3:8.2f} us/step", flush=True)
            del groups
