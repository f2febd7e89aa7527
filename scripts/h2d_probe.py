"""Raw pinned H2D / D2H rate of the box (the e2e floor of nldsc_ld_calculate_packed is packed bytes / this)."""
import torch, time
n = 250_000_000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); h.random_(0, 255)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
whole = t(lambda: d.copy_(h, non_blocking=True))
def pieces():
    for o in range(0, n, 8 << 20): d[o:o + (8 << 20)].copy_(h[o:o + (8 << 20)], non_blocking=True)
pc = t(pieces)
back = t(lambda: h.copy_(d, non_blocking=True))
print(f"H2D whole {n / whole / 1e9:.1f} GB/s ({whole * 1e3:.2f} ms), 8MB pieces {n / pc / 1e9:.1f} GB/s, D2H {n / back / 1e9:.1f} GB/s", flush=True)
