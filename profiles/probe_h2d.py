"""PCIe ceiling of the box: one 2 GB host-to-device copy from pinned memory vs 192 copies (what a step of\nbench.py ships).  python profiles/probe_h2d.py"""
import torch, time
n = 1996800000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def run(chunks):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for it in range(5):
        e0.record()
        step = n // chunks
        for c in range(chunks):
            d[c*step:(c+1)*step].copy_(h[c*step:(c+1)*step], non_blocking=True)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return n / best / 1e6
print("H2D pinned, 1 copy :", round(run(1), 1), "GB/s")
print("H2D pinned, 192 copies:", round(run(192), 1), "GB/s")
