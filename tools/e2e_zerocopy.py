"""Experiment: the persistent kernel reading / writing PINNED HOST memory directly (UVA), no staging copies."""
import ctypes as C, os, sys, time
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench
from maniskill_b200 import _lib as L

B = 1 << 20
g = bench.Group("panda", B, 2022, "delta", True)
kin = g.ctrl.kinematics
d_q = g.art.px.cuda_articulation_qpos.torch()
h_q = d_q.cpu().pin_memory(); h_a = g.action.cpu().pin_memory()
h_o = torch.empty((B, 7), dtype=torch.float32).pin_memory()
h_ref = torch.empty((B, 7), dtype=torch.float32).pin_memory()
norm = ((-0.1,) * 3, (0.1,) * 3, -0.1)
kin.delta_step_host(h_q, h_a, h_ref, normalize=norm)
cfg = kin.step_cfg(B, 9, 9, 7, 0, L.F_USE_DELTA | L.F_NO_DRIVE | L.F_NORMALIZE, pos_low=norm[0], pos_high=norm[1], rot_scale=norm[2])
st = torch.cuda.current_stream().cuda_stream
def zc():
    L.check(kin._lib.msik_delta_step(kin._chain, C.byref(cfg), h_q.data_ptr(), None, h_a.data_ptr(), None, h_o.data_ptr(), None, st))
    torch.cuda.synchronize()
for _ in range(3): zc()
print("bitwise equal to the staged pipeline:", torch.equal(h_o, h_ref))
t0 = time.perf_counter(); K = 20
for _ in range(K): zc()
dt = (time.perf_counter() - t0) / K
print(f"zero-copy: {dt * 1e3:.3f} ms  {B / dt / 1e8:.2f}e8 envs/s  host->device {B * 64 / dt / 1e9:.1f} GB/s")
t0 = time.perf_counter()
for _ in range(K): kin.delta_step_host(h_q, h_a, h_ref, normalize=norm)
dt = (time.perf_counter() - t0) / K
print(f"staged   : {dt * 1e3:.3f} ms  {B / dt / 1e8:.2f}e8 envs/s")
