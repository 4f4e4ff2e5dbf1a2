"""Pinned D2H bandwidth of a 4096^2 RGBA32F frame on this box (the floor of the e2e frame time)."""
import time
import torch

x = torch.zeros((4096, 4096, 4), dtype=torch.float32, device="cuda")
h = torch.zeros((4096, 4096, 4), dtype=torch.float32).pin_memory()
for n in (1, 8):
    parts = x.view(n, -1), h.view(n, -1)
    for _ in range(3):
        for i in range(n):
            parts[1][i].copy_(parts[0][i], non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        for i in range(n):
            parts[1][i].copy_(parts[0][i], non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    print(f"D2H 268 MB in {n} piece(s): {dt * 1e3:.2f} ms = {x.numel() * 4 / dt / 1e9:.1f} GB/s")
