"""H2D of 64 MiB (+ concurrent D2H of 28 MiB) from pinned host memory split over 1 / 2 / 4 streams: does more than one copy
stream per direction raise the PCIe throughput of this box?  (dev tool; the e2e leg of bench.py uses one stream per direction.)"""
import torch

H, D = 64 << 20, 28 << 20
h_in = torch.empty(H, dtype=torch.uint8).pin_memory()
h_out = torch.empty(D, dtype=torch.uint8).pin_memory()
d_in = torch.empty(H, dtype=torch.uint8, device="cuda")
d_out = torch.empty(D, dtype=torch.uint8, device="cuda")
for n in (1, 2, 4):
    for both in (False, True):
        ss = [torch.cuda.Stream() for _ in range(n)]
        so = torch.cuda.Stream()
        best = 1e9
        for rep in range(12):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for s in ss + [so]:
                s.wait_event(e0)
            per = H // n
            for i, s in enumerate(ss):
                with torch.cuda.stream(s):
                    d_in[i * per:(i + 1) * per].copy_(h_in[i * per:(i + 1) * per], non_blocking=True)
            if both:
                with torch.cuda.stream(so):
                    h_out.copy_(d_out, non_blocking=True)
            for s in ss + [so]:
                torch.cuda.current_stream().wait_stream(s)
            e1.record()
            torch.cuda.synchronize()
            if rep >= 2:
                best = min(best, e0.elapsed_time(e1))
        print(f"H2D 64 MiB over {n} stream(s){' + D2H 28 MiB' if both else '':14s}: {best:.3f} ms  ({H / best / 1e6:.1f} GB/s H2D)", flush=True)
