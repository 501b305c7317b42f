"""Quick on-GPU sanity check of the kernels against the fp64 oracle (dev tool)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from maniskill_b200.kinematics import Kinematics
from maniskill_b200.shim import FakeArticulation
from maniskill_b200.structs.pose import Pose
from oracle import oracle as O

A = os.path.join(os.path.dirname(__file__), "..", "maniskill_b200", "assets", "robots")

def run(name, ee, ctrl_names, B=4096, seed=0):
    urdf = os.path.join(A, name)
    art = FakeArticulation(urdf, B, ee_link_names=[ee])
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in ctrl_names], dtype=torch.int32, device="cuda")
    kin = Kinematics(urdf, ee, art, idx)
    fr = O.load_chain(urdf, ee)
    lim = O.joint_limits(fr); lim = np.clip(lim, -3.0, 3.0)
    rng = np.random.default_rng(seed)
    lo, hi = lim[:, 0] + 0.1 * (lim[:, 1] - lim[:, 0]), lim[:, 1] - 0.1 * (lim[:, 1] - lim[:, 0])
    n = len(lim)
    qa = rng.uniform(lo, hi, (B, n))
    qpos = np.zeros((B, art.max_dof), np.float32)
    anc = kin.active_ancestor_joint_idxs
    qpos[:, anc] = qa
    qa = qpos[:, anc].astype(np.float64)
    q0 = torch.from_numpy(qpos).cuda()
    print(name, "kernel:", kin.kernel_name, "n=", n, "ctrl=", kin.n_ctrl)
    # FK + J
    pose, jac = kin.forward_kinematics(q0, return_jacobian=True)
    J, R, p = O.jacobian(fr, qa)
    print("  fk pos err", np.abs(pose.p.cpu().numpy() - p).max(), "jac err", np.abs(jac.cpu().numpy() - J).max())
    qo = O.matrix_to_quaternion(R)
    qe = pose.q.cpu().numpy()
    print("  fk quat err", np.minimum(np.abs(qe - qo).max(1), np.abs(qe + qo).max(1)).max())
    # path A
    a = rng.uniform(-1, 1, (B, 6)) * 0.1
    mask = np.zeros(n, bool); mask[kin.controlled_joints_idx_in_qmask] = True
    ref = O.ik_delta_step(fr, qa, a, col_mask=mask)
    out = kin.compute_ik(None, q0, action=torch.from_numpy(a.astype(np.float32)).cuda(), use_delta_ik_solver=True)
    err = np.abs(out.cpu().numpy() - ref).max(1)
    cond = np.linalg.cond(J[:, :, mask])
    for lo_c, hi_c in [(0, 50), (50, 200), (200, 1e3), (1e3, 1e9)]:
        m = (cond >= lo_c) & (cond < hi_c)
        if m.any(): print(f"  A cond[{lo_c},{hi_c}): n={m.sum()} max err {err[m].max():.2e} median {np.median(err[m]):.2e}")
    # path A pos only
    ref3 = O.ik_delta_step(fr, qa, a[:, :3], pos_only=True, col_mask=mask)
    out3 = kin.compute_ik(None, q0, pos_only=True, action=torch.from_numpy(a[:, :3].astype(np.float32)).cuda(), use_delta_ik_solver=True)
    print("  A pos-only max err", np.abs(out3.cpu().numpy() - ref3).max())
    # path B: target = FK of perturbed q
    Bb = min(B, 1024)
    q2 = qa[:Bb].copy(); q2[:, mask] += rng.uniform(-0.15, 0.15, (Bb, mask.sum()))
    _, R2, p2 = O.jacobian(fr, q2)
    tp = np.concatenate([p2, O.matrix_to_quaternion(R2)], 1)
    refB, its = O.ik_iterative(fr, qa[:Bb], tp, col_mask=mask, return_info=True)
    outB = kin.compute_ik(Pose.create(torch.from_numpy(tp.astype(np.float32)).cuda()), q0[:Bb])
    eB = np.abs(outB.cpu().numpy() - refB).max(1)
    print(f"  B iters mean {its.mean():.1f} max {its.max()}  err max {eB.max():.2e} median {np.median(eB):.2e} frac>1e-5 {(eB>1e-5).mean():.4f}")

if __name__ == "__main__":
    run("panda_v2_kin.urdf", "panda_hand_tcp", [f"panda_joint{i}" for i in range(1, 8)])
    run("fetch_kin.urdf", "gripper_link", ["shoulder_pan_joint","shoulder_lift_joint","upperarm_roll_joint","elbow_flex_joint","forearm_roll_joint","wrist_flex_joint","wrist_roll_joint"])
    run("xarm7_ability_right_hand_kin.urdf", "base", [f"joint{i}" for i in range(1, 8)])
    os.environ["MSIK_FORCE_GENERIC"] = "1"
    run("panda_v2_kin.urdf", "panda_hand_tcp", [f"panda_joint{i}" for i in range(1, 8)])
    run("fetch_kin.urdf", "gripper_link", ["shoulder_pan_joint","shoulder_lift_joint","upperarm_roll_joint","elbow_flex_joint","forearm_roll_joint","wrist_flex_joint","wrist_roll_joint"])
