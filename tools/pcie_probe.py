"""Host<->device copy rates of the box (pinned memory), alone and concurrent: the bound of bench.py's e2e number."""
import time
import torch

def rate(fn, nbytes, reps=10):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return nbytes * reps / (time.perf_counter() - t0) / 1e9

h_in = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
h_out = torch.empty(28 << 20, dtype=torch.uint8).pin_memory()
d_in = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
d_out = torch.empty(28 << 20, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)
def both():
    h2d(); d2h()

print("H2D alone  %.1f GB/s" % rate(h2d, 64 << 20))
print("D2H alone  %.1f GB/s" % rate(d2h, 28 << 20))
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    both()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 10
print("concurrent 64 MiB in + 28 MiB out: %.3f ms per pair  (H2D %.1f GB/s while D2H runs)" % (dt * 1e3, (64 << 20) / dt / 1e9))
for mb in (2, 8, 16):
    n = mb << 20
    def chunks():
        with torch.cuda.stream(s1):
            for o in range(0, 64 << 20, n):
                d_in[o:o + n].copy_(h_in[o:o + n], non_blocking=True)
    print("H2D in %2d MiB pieces %.1f GB/s" % (mb, rate(chunks, 64 << 20)))
