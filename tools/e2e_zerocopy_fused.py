"""Experiment: the fused Panda step (arm + mimic gripper, whole target rows) with ALL its buffers in pinned host memory -- the
persistent kernel's bulk copies (cp.async.bulk, 0.9 - 1.2 KB each) then read qpos / action rows and write target rows straight
over PCIe, no staging copies and no pipeline tail.  How close does that get to the link?  (dev tool)"""
import ctypes as C, os, sys, time
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench
from maniskill_b200 import _lib as L

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
g = bench.Group("panda", B, 2022, "delta", True)
kin = g.ctrl.kinematics
lib = kin._lib
d_q = g.art.px.cuda_articulation_qpos.torch()
d_t = g.art.px.cuda_articulation_target_qpos.torch()
h_q = d_q.cpu().pin_memory(); h_a = g.action.cpu().pin_memory()
h_t = torch.zeros((B, 9), dtype=torch.float32).pin_memory()
norm = ((-0.1,) * 3, (0.1,) * 3, -0.1)
cfg = kin.step_cfg(B, 9, 9, 7, 0, L.F_USE_DELTA | L.F_NORMALIZE, pos_low=norm[0], pos_high=norm[1], rot_scale=norm[2])
aux = L.AuxCfg()
aux.n_cols = 2
for i, col in enumerate((7, 8)):
    aux.col[i] = col; aux.action_col[i] = 6; aux.flags[i] = L.AUX_NORMALIZE
    aux.mid[i] = 0.5 * (0.04 + -0.01); aux.half[i] = 0.5 * (0.04 - -0.01)
st = torch.cuda.current_stream().cuda_stream
def run(q, a, t):
    L.check(lib.msik_delta_step_fused(kin._chain, C.byref(cfg), C.byref(aux), q.data_ptr(), None, a.data_ptr(), t.data_ptr(), None, None, st))
run(d_q, g.action, d_t); torch.cuda.synchronize()
want = d_t.clone()  # the same call on device-resident buffers
for _ in range(3): run(h_q, h_a, h_t)
torch.cuda.synchronize()
print("host (zero-copy) buffers give the same target rows as device buffers:", torch.equal(h_t.cuda(), want))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 20
e0.record()
for _ in range(K): run(h_q, h_a, h_t)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
print(f"zero-copy fused step: {ms:.3f} ms per {B} envs = {B / ms / 1e5:.2f}e8 envs/s; host->device {B * 64 / ms / 1e6:.1f} GB/s + device->host {B * 36 / ms / 1e6:.1f} GB/s")
