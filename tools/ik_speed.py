"""Path B (iterative IK, BASELINE configs[2]: Panda pd_ee_pose, absolute targets) time per step at several env counts,
through the controller (bench.py's Group), L2 flushed between steps, plus the iteration statistics (dev tool).
MSIK_LIB=<variant build> selects another library for A/B runs; IK_REG / IK_GUARD override regularization / guard_pivot."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench as BN

sizes = [int(x) for x in (sys.argv[1:] or ["4096", "16384", "65536", "262144"])]
for B in sizes:
    g = BN.Group("panda", B, 0, "pd_ee_pose")
    kin = g.ctrl.kinematics
    if "IK_REG" in os.environ: kin.regularization = float(os.environ["IK_REG"])
    if "IK_GUARD" in os.environ: kin.guard_pivot = float(os.environ["IK_GUARD"])
    if "IK_MAP" in os.environ: kin.ik_mapping = os.environ["IK_MAP"]
    for _ in range(3): g.step()
    flush = torch.empty(2 * BN.L2_BYTES // 4, dtype=torch.float32, device="cuda")
    K = 30
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for a, b in ev:
        flush.zero_(); a.record(); g.step(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
    q = g.art.px.cuda_articulation_qpos.torch()
    tp = g.ctrl._target_pose
    _, info = kin.compute_ik(tp, q, return_info=True)
    it = info["iterations"]
    parked = int((it < 0).sum())
    it = it.abs().float()
    print(f"B={B:7d}  median {ts[K // 2]:8.2f} us  min {ts[0]:8.2f} us   {B / ts[K // 2] * 1e6:.3e} envs*steps/s   "
          f"iters mean {it.mean():.2f} max {int(it.max())}  parked {parked}  err max pos {info['error'][:, 0].max():.2e} rot {info['error'][:, 1].max():.2e}", flush=True)
    del g
