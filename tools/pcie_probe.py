"""Raw pinned host<->device copy bandwidth of this box (context for the end-to-end number)."""
import sys, torch
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return n * reps / (a.elapsed_time(b) * 1e-3) / 1e9
print(f"H2D {t(lambda: d.copy_(h, non_blocking=True)):.1f} GB/s  D2H {t(lambda: h.copy_(d, non_blocking=True)):.1f} GB/s", file=sys.stderr)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
h2 = torch.empty(n, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize()
import time
both(); torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): both()
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"duplex: {n * 5 / dt / 1e9:.1f} GB/s each way", file=sys.stderr)
