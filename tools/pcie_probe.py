"""H2D / D2H bandwidth of pinned copies on this box (context for the e2e number)."""
import time
import torch
n = 512 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, src, dst in (("h2d", h, d), ("d2h", d, h)):
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(5):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / 5
    print(f"{name}: {n / dt / 1e9:.1f} GB/s")
