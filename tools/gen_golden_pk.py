#!/usr/bin/env python3
"""Pin-on-arrival for the two absent third-party solvers (missing #3).

The arithmetic of paths B and C lives in ``pytorch_kinematics==0.7.5`` (setup.py:46) and ``sapien`` (setup.py:50-51),
neither installable in this container.  The day one of them imports -- from ``baseline/_ref`` or site-packages -- this
script dumps what the oracle could only restate so far, and tests/test_pk_golden.py (skipped with that reason today)
becomes binding:

  pytorch_kinematics:  chain.jacobian(q) and forward kinematics for the shipped chains (call sites kinematics.py:106-109,
                       157), the REAL defaults of PseudoInverseIK as constructed at kinematics.py:110-117, and
                       solve() results (solutions[:, 0, :], iterations, convergence flags) as called at :145-155
  sapien:              PinocchioModel.compute_inverse_kinematics results as called at kinematics.py:174-180

Usage:  python tools/gen_golden_pk.py     (exit 0 and a one-line note when neither package is importable)
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
OUT = os.path.join(ROOT, "tests", "golden", "pk_golden.npz")
ASSETS = os.path.join(ROOT, "maniskill_b200", "assets", "robots")
CHAINS = {"panda": ("panda_v2_kin.urdf", "panda_hand_tcp"), "fetch": ("fetch_kin.urdf", "gripper_link")}
PANDA_REST = np.array([0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4])


def dump_pk(out: dict) -> bool:
    try:
        import pytorch_kinematics as pk
        import torch
    except Exception as e:  # noqa: BLE001
        print(f"pytorch_kinematics not importable ({type(e).__name__}: {e}): paths A-jacobian / B stay restated")
        return False
    out["pk_version"] = np.array(getattr(pk, "__version__", "?"))
    rng = np.random.default_rng(2022)
    for name, (urdf, ee) in CHAINS.items():
        with open(os.path.join(ASSETS, urdf), "rb") as f:
            chain = pk.build_serial_chain_from_urdf(f.read(), end_link_name=ee).to(dtype=torch.float32)
        lim = np.stack([np.asarray(x, dtype=np.float64) for x in chain.get_joint_limits()], 1)  # (n, 2)
        lim = np.clip(lim, -3.0, 3.0)
        n = lim.shape[0]
        q = rng.uniform(lim[:, 0] * 0.8, lim[:, 1] * 0.8, (256, n)).astype(np.float32)
        J, T = chain.jacobian(torch.from_numpy(q), ret_eef_pose=True)
        out[f"{name}_q"] = q
        out[f"{name}_jacobian"] = J.numpy()
        out[f"{name}_eef_matrix"] = T.numpy() if hasattr(T, "numpy") else T.get_matrix().numpy()
        if name != "panda":
            continue
        # PseudoInverseIK exactly as the reference constructs it (kinematics.py:110-117)
        lim_t = torch.tensor(np.stack([np.asarray(x, dtype=np.float32) for x in chain.get_joint_limits()], 1))
        pik = pk.PseudoInverseIK(chain, joint_limits=lim_t, early_stopping_any_converged=True, max_iterations=200, num_retries=1)
        for attr in ("lr", "regularlization", "regularization", "pos_tolerance", "rot_tolerance", "max_iterations",
                     "early_stopping_no_improvement", "early_stopping_no_improvement_patience", "line_search", "optimizer_method"):
            if hasattr(pik, attr):
                v = getattr(pik, attr)
                out[f"pik_default_{attr}"] = np.array(str(v) if not isinstance(v, (int, float, bool)) else v)
        q0 = (PANDA_REST[None] + 0.02 * rng.normal(size=(128, 7))).astype(np.float32)
        q2 = q0 + (0.3 * (rng.random((128, 7)) - 0.5)).astype(np.float32)
        Tt = chain.forward_kinematics(torch.from_numpy(q2))
        M = Tt.get_matrix()
        pos = M[:, :3, 3]
        rot = pk.matrix_to_quaternion(M[:, :3, :3])
        pik.initial_config = torch.from_numpy(q0)
        sol = pik.solve(pk.Transform3d(pos=pos, rot=rot))
        out["pik_q0"] = q0
        out["pik_target_pos"] = pos.numpy()
        out["pik_target_quat_wxyz"] = rot.numpy()
        out["pik_solutions"] = sol.solutions[:, 0, :].numpy()
        for attr in ("iterations", "converged", "err_pos", "err_rot"):
            if hasattr(sol, attr):
                v = getattr(sol, attr)
                out[f"pik_{attr}"] = np.asarray(v.numpy() if hasattr(v, "numpy") else v)
    return True


def dump_sapien(out: dict) -> bool:
    try:
        import sapien
        from sapien.wrapper.pinocchio_model import PinocchioModel  # noqa: F401
    except Exception as e:  # noqa: BLE001
        print(f"sapien / PinocchioModel not importable ({type(e).__name__}: {e}): path C stays restated")
        return False
    # needs a loaded articulation (physx) -- recorded here for the maintainer who has SAPIEN: build the Panda from
    # mani_skill/assets/robots/panda/panda_v2.urdf, pmodel = articulation.create_pinocchio_model(), then for seeded
    # (q0, target) pairs store pmodel.compute_inverse_kinematics(ee_link_idx, pose, initial_qpos=q0, active_qmask=mask,
    # max_iterations=100) -> (result, success, error) under clik_q0 / clik_target / clik_result / clik_success.
    out["sapien_version"] = np.array(getattr(sapien, "__version__", "?"))
    print("sapien imports: extend dump_sapien() on a machine where physx can load an articulation")
    return True


def main() -> int:
    out: dict = {}
    ok = dump_pk(out)
    ok = dump_sapien(out) or ok
    if not ok:
        print("nothing to pin: tests/golden/pk_golden.npz not written (tests/test_pk_golden.py stays skipped)")
        return 0
    np.savez_compressed(OUT, **out)
    print(f"wrote {os.path.relpath(OUT)} with {sorted(out)}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
