"""D2H / H2D bandwidth of this box with pinned memory (context for the e2e leg of bench.py)."""
import torch, time
n = 310 * 1024 * 1024
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h = torch.empty(n, dtype=torch.uint8).pin_memory()
for name, f in (("D2H", lambda: h.copy_(d, non_blocking=True)), ("H2D", lambda: d.copy_(h, non_blocking=True))):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): f()
    e1.record(); torch.cuda.synchronize()
    print(f"{name} {5 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9:.1f} GB/s (310 MiB pinned)")
for mb in (1, 4, 25, 39):
    m = mb * 1024 * 1024
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20): h[:m].copy_(d[:m], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(f"D2H {mb} MiB pieces: {20 * m / (e0.elapsed_time(e1) * 1e-3) / 1e9:.1f} GB/s")
