import torch, time
n = 192; sz = 10_400_000
host = torch.empty(n * sz, dtype=torch.uint8, pin_memory=True)
dev = torch.empty(n * sz, dtype=torch.uint8, device="cuda")
s = torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.time() - t0) / reps
def many():
    with torch.cuda.stream(s):
        for i in range(n):
            dev[i*sz:(i+1)*sz].copy_(host[i*sz:(i+1)*sz], non_blocking=True)
def one():
    with torch.cuda.stream(s):
        dev.copy_(host, non_blocking=True)
def four():
    q = n * sz // 4
    with torch.cuda.stream(s):
        for i in range(4):
            dev[i*q:(i+1)*q].copy_(host[i*q:(i+1)*q], non_blocking=True)
for name, fn in (("192 copies", many), ("1 copy", one), ("4 copies", four)):
    dt = t(fn)
    print(f"{name}: {dt*1e3:.2f} ms = {n*sz/dt/1e9:.1f} GB/s")
