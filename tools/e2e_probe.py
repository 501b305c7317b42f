"""End-to-end (host buffers) timing of Kinematics.delta_step_host vs chunk size (dev tool)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench

B = 1 << 20
g = bench.Group("panda", B, 2022, "delta", True)
kin = g.ctrl.kinematics
h_q = g.art.px.cuda_articulation_qpos.torch().cpu().pin_memory()
h_a = g.action.cpu().pin_memory()
h_o = torch.empty((B, 7), dtype=torch.float32).pin_memory()
norm = ((-0.1,) * 3, (0.1,) * 3, -0.1)
for chunk in [1 << 15, 1 << 16, 1 << 17, 3 << 16, 1 << 18, 1 << 19]:
    for _ in range(3):
        kin.delta_step_host(h_q, h_a, h_o, normalize=norm, chunk_envs=chunk)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    K = 20
    for _ in range(K):
        kin.delta_step_host(h_q, h_a, h_o, normalize=norm, chunk_envs=chunk)
    dt = (time.perf_counter() - t0) / K
    print(f"chunk {chunk:8d}: {dt * 1e3:.3f} ms  {B / dt / 1e8:.2f}e8 envs/s  h2d {B * 64 / dt / 1e9:.1f} GB/s", flush=True)
