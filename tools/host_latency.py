"""Where the host time of one small-batch controller step goes (4096 Panda envs, stock fused dict, steps back to back on
L2-resident buffers): public API -> program closure -> bare foreign call -> CUDA-graph replay (dev tool)."""
import ctypes as C, os, sys, time
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench as BN

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
g = BN.Group("panda", B, 0, "pd_ee_delta_pose", True)
cc, a = g.combined, g.action
N = 20000

def t(fn, n=N):
    for _ in range(200): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6

print(f"B={B}")
print(f"CombinedController.set_action (public API)   {t(lambda: cc.set_action(a)):6.2f} us")
prog = cc._program()
print(f"program step closure                          {t(lambda: prog(a)):6.2f} us")
arm = cc.controllers["arm"]
ptr = a.data_ptr(); st = torch.cuda.current_stream().cuda_stream
h = arm._step_handles[0]
f2 = arm._lib.msik_step_launch
print(f"bare ctypes msik_step_launch                  {t(lambda: f2(h, ptr, st)):6.2f} us")
a2 = [a.clone() for _ in range(4)]
k = [0]
def fresh():
    k[0] += 1
    cc.set_action(a2[k[0] & 3])
print(f"set_action, a different action tensor per step {t(fresh):6.2f} us")
rep = cc.capture(a)
print(f"capture() replay (1 kernel node)              {t(rep):6.2f} us")
x = torch.zeros(1, device="cuda")
print(f"torch x.add_(1) (smallest torch launch)       {t(lambda: x.add_(1)):6.2f} us")
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    for _ in range(16): cc.set_action(a)
print(f"16 steps per graph replay, per step           {t(gr.replay, 2000) / 16:6.2f} us")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
print(f"device time per step inside that graph        {e0.elapsed_time(e1) / 16 * 1e3:6.2f} us")
