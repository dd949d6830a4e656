"""Pinned host->device copy bandwidth of this box (the e2e leg's ceiling): python tools/h2d_bw.py"""
import torch
n = 88080384
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
o = torch.empty(1835008, dtype=torch.uint8).pin_memory()
od = torch.empty(1835008, dtype=torch.uint8, device="cuda")
for _ in range(3): d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): d.copy_(h, non_blocking=True)
e1.record(); e1.synchronize()
t = e0.elapsed_time(e1) / 20 * 1e-3
print(f"H2D {n/1e6:.1f} MB: {t*1e3:.3f} ms -> {n/t/1e9:.1f} GB/s  (= {8/t:.0f} pairs/s ceiling for batch-8 fp32 1024x448 input)")
s2 = torch.cuda.Stream()
e0.record()
for _ in range(20):
    d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): o.copy_(od, non_blocking=True)
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 20 * 1e-3
print(f"H2D with concurrent D2H: {t*1e3:.3f} ms -> {n/t/1e9:.1f} GB/s")
