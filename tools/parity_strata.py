"""Joint-target error of path A against the fp64 oracle per stratum of cond(J), three columns:
  ours        libmsik (reduced / dense / table-driven solve, fp32)
  ref-fp32    the reference's own GPU arithmetic restated in eager torch fp32 (baseline/eager_ref.py:
              per-frame Jacobian + torch.linalg.pinv, kinematics.py:157,171) -- pytorch_kinematics is absent
  both against oracle.ik_delta_step (fp64 numpy pinv with torch's fp32 cut-off).
Run on a B200:  python tools/parity_strata.py > profiles/rNN_parity_strata.txt"""
import os, sys
import numpy as np
import torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import FETCH, PANDA, REALMAN_R, XARM7
from helpers import make_art, sample_q
from baseline.eager_ref import EagerChain
from maniskill_b200.kinematics import Kinematics
from oracle import oracle as O

EDGES = ((0, 50), (50, 200), (200, 1e3), (1e3, 1e12))


def strata(robot, B=8192, generic=False):
    if generic: os.environ["MSIK_FORCE_GENERIC"] = "1"
    art = make_art(robot, B)
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in robot[2]], dtype=torch.int32, device="cuda")
    kin = Kinematics(robot[0], robot[1], art, idx)
    os.environ.pop("MSIK_FORCE_GENERIC", None)
    fr = O.load_chain(robot[0], robot[1])
    mask = np.zeros(kin.chain_table.n, bool); mask[kin.controlled_joints_idx_in_qmask] = True
    q = sample_q(fr, B, seed=2)
    rows = np.random.default_rng(1).uniform(0, 0.04, (B, art.max_dof)).astype(np.float32)
    rows[:, kin.active_ancestor_joint_idxs] = q.astype(np.float32)
    qa = rows[:, kin.active_ancestor_joint_idxs].astype(np.float64)
    a = np.random.default_rng(3).uniform(-0.1, 0.1, (B, 6)).astype(np.float32)
    ref = O.ik_delta_step(fr, qa, a.astype(np.float64), col_mask=mask)
    q0 = torch.from_numpy(rows).cuda()
    at = torch.from_numpy(a).cuda()
    ours = kin.compute_ik(None, q0, action=at, use_delta_ik_solver=True).cpu().numpy()
    # reference arithmetic in fp32 (masked columns = the intended J[:, :, qmask])
    ch = EagerChain(robot[0], robot[1], "cuda")
    qt = q0[:, kin.active_ancestor_joint_idxs]
    J32 = ch.jacobian(qt)[:, :, kin.controlled_joints_idx_in_qmask]
    eager = (qt[:, kin.controlled_joints_idx_in_qmask] + (torch.linalg.pinv(J32) @ at.unsqueeze(-1)).squeeze(-1)).cpu().numpy()
    cond = np.linalg.cond(O.jacobian(fr, qa)[0][:, :, mask])
    eo, er = np.abs(ours - ref).max(1), np.abs(eager - ref).max(1)
    for lo, hi in EDGES:
        m = (cond >= lo) & (cond < hi)
        if m.any():
            print(f"{kin.kernel_name:18s} {lo:>6g}-{hi:<8g} {int(m.sum()):6d}   {eo[m].max():10.2e} {eo[m].mean():10.2e}   {er[m].max():10.2e} {er[m].mean():10.2e}   "
                  f"{'ours <= ref' if eo[m].mean() <= er[m].mean() else 'ours > ref'}", flush=True)


if __name__ == "__main__":
    print("# tools/parity_strata.py on B200: |q_target - oracle(fp64 pinv)| per stratum of cond(J), 8192 random configurations inside the joint")
    print("# limits per robot, actions U(-0.1, 0.1)^6; ours = libmsik kernel named in column 1, ref-fp32 = eager torch fp32 restatement of the")
    print("# reference GPU path (per-frame Jacobian + torch.linalg.pinv); (max, mean) in rad")
    print(f"{'kernel':18s} {'cond(J)':>15s} {'n':>6s}   {'ours max':>10s} {'ours mean':>10s}   {'ref32 max':>10s} {'ref32 mean':>10s}")
    for robot in (PANDA, FETCH, XARM7, REALMAN_R):
        strata(robot)
    strata(PANDA, generic=True)
    strata(FETCH, generic=True)
