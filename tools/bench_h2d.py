"""Pinned host -> device copy bandwidth of the box (the ceiling of bench.py's end-to-end number)."""
import torch
n = 3774873600  # one bench step of frames
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda:0")
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    d.copy_(h, non_blocking=True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"H2D pinned {n / 1e9:.2f} GB in {ms:.1f} ms = {n / ms / 1e6:.1f} GB/s -> at most {4096 / (ms * 1e-3):.0f} frames/s of 640x480 BGR")
