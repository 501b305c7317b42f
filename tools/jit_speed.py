"""Table-driven vs run-time specialised kernels on a chain the library has no built-in code for (dev tool)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from helpers import make_art
from maniskill_b200.kinematics import Kinematics

URDF = os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "toy6_kin.urdf")
TOY = (URDF, "tool", [f"j{i}" for i in range(1, 7)])
B = 1 << 20
for spec in (False, True):
    art = make_art(TOY, 8)
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in TOY[2]], dtype=torch.int32, device="cuda")
    if not spec: os.environ["MSIK_FORCE_GENERIC"] = "1"
    kin = Kinematics(TOY[0], TOY[1], art, idx, specialise=spec)
    os.environ.pop("MSIK_FORCE_GENERIC", None)
    qs = [torch.rand((B, 6), device="cuda") * 2 - 1 for _ in range(6)]   # 6 x 25 MB > L2 with actions and outputs
    as_ = [(torch.rand((B, 6), device="cuda") * 2 - 1) * 0.05 for _ in range(6)]
    for i in range(6): kin.compute_ik(None, qs[i], action=as_[i], use_delta_ik_solver=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    K = 60
    for i in range(K): kin.compute_ik(None, qs[i % 6], action=as_[i % 6], use_delta_ik_solver=True)
    e1.record(); torch.cuda.synchronize()
    print(f"{kin.kernel_name:20s} {e0.elapsed_time(e1) / K * 1e3:8.2f} us per 2^20 envs (path A, NO_DRIVE, incl. output allocation)")
