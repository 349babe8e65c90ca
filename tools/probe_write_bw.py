"""Pure-write HBM bandwidth next to the copy figure of MEASURED_PEAKS.json (development aid)."""
import torch
x = torch.empty(2 * 1024 ** 3 // 8, dtype=torch.float64, device="cuda")
y = torch.empty_like(x)
def timeit(f, n=10):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(n):
        a.record(); f(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
t = timeit(lambda: x.fill_(1.5))
print(f"fill 2 GiB: {t:.3f} ms = {x.numel() * 8 / t / 1e6:.0f} GB/s written")
t = timeit(lambda: y.copy_(x))
print(f"copy 2 GiB: {t:.3f} ms = {2 * x.numel() * 8 / t / 1e6:.0f} GB/s read+write")
t = timeit(lambda: torch.cuda.memset if False else x.zero_())
print(f"zero 2 GiB: {t:.3f} ms = {x.numel() * 8 / t / 1e6:.0f} GB/s written")
s = timeit(lambda: x.sum())
print(f"sum 2 GiB: {s:.3f} ms = {x.numel() * 8 / s / 1e6:.0f} GB/s read")
