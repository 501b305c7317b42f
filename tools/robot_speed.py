"""Path-A time per 2^20 envs for every robot with generated chain code: Kinematics.compute_ik(use_delta_ik_solver=True)
(NO_DRIVE, output tensor allocated per call), generated kernel vs the table-driven one (dev tool)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import FETCH, PANDA, REALMAN_R, XARM7
from helpers import make_art, sample_q
from maniskill_b200.kinematics import Kinematics
from oracle import oracle as O

B = 1 << 20
for robot in (PANDA, FETCH, XARM7, REALMAN_R):
    for generic in (False, True):
        if generic: os.environ["MSIK_FORCE_GENERIC"] = "1"
        art = make_art(robot, 8)
        act = [j.name for j in art.active_joints]
        idx = torch.tensor([act.index(n) for n in robot[2]], dtype=torch.int32, device="cuda")
        kin = Kinematics(robot[0], robot[1], art, idx)
        os.environ.pop("MSIK_FORCE_GENERIC", None)
        fr = O.load_chain(robot[0], robot[1])
        nd = art.max_dof
        qa = torch.from_numpy(sample_q(fr, 4096, seed=1).astype(np.float32)).cuda()
        qs = []
        for g in range(6):
            rows = torch.rand((B, nd), device="cuda") * 0.04
            rows[:, kin.active_ancestor_joint_idxs] = qa[torch.randint(0, 4096, (B,), device="cuda")]
            qs.append(rows)
        as_ = [(torch.rand((B, 6), device="cuda") * 2 - 1) * 0.05 for _ in range(6)]
        for i in range(6): kin.compute_ik(None, qs[i], action=as_[i], use_delta_ik_solver=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        K = 60
        for i in range(K): kin.compute_ik(None, qs[i % 6], action=as_[i % 6], use_delta_ik_solver=True)
        e1.record(); torch.cuda.synchronize()
        print(f"{os.path.basename(robot[0]):40s} {kin.kernel_name:18s} row {nd:2d} floats  {e0.elapsed_time(e1) / K * 1e3:8.2f} us per 2^20 envs", flush=True)
        del qs, as_
