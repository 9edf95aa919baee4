"""Development aid: pinned host <-> device copy bandwidth of this box (the PCIe roofline for e2e)."""
import torch, time
n = 2 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 2, dtype=torch.uint8, pin_memory=True)
d2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return best
a = t(lambda: d.copy_(h, non_blocking=True))
b = t(lambda: h.copy_(d, non_blocking=True))
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
c = t(both)
print("H2D %.1f GB/s  D2H %.1f GB/s  duplex: H2D %d MB + D2H %d MB in %.1f ms -> H2D-equivalent %.1f GB/s" % (
    n / a / 1e9, n / b / 1e9, n >> 20, n >> 21, c * 1e3, n / c / 1e9))
