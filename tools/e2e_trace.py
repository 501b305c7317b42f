import os, sys, time, torch
sys.path.insert(0, "/root/repo")
import bench
B = 1 << 20
g = bench.Group("panda", B, 2022, "delta", True)
kin = g.ctrl.kinematics
h_q = g.art.px.cuda_articulation_qpos.torch().cpu().pin_memory(); h_a = g.action.cpu().pin_memory()
h_o = torch.empty((B, 7), dtype=torch.float32).pin_memory()
norm = ((-0.1,) * 3, (0.1,) * 3, -0.1)
for chunk in (1 << 17, 1 << 18):
    for _ in range(3):
        t0 = time.perf_counter(); kin.delta_step_host(h_q, h_a, h_o, normalize=norm, chunk_envs=chunk); print("wall ms", (time.perf_counter() - t0) * 1e3, file=sys.stderr)
