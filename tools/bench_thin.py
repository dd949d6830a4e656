"""Thin pyramid layers: tcgen05 conv vs the thin-layer kernel (GPU durations via ncu: run under
ncu --metrics gpu__time_duration.sum -k regex:"conv_tc|conv_thin" ...).  python tools/bench_thin.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
d = torch.device("cuda")
cases = [("conv1aa/b", 16, 16, 1, 16, 224, 512), ("conv2a", 16, 32, 2, 16, 224, 512), ("conv2aa/b", 32, 32, 1, 16, 112, 256),
         ("conv3aa/b", 64, 64, 1, 16, 56, 128)]
for name, cin, cout, stride, N, H, W in cases:
    x = torch.randn(N, H, W, cin, device=d).bfloat16()
    w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
    b = torch.zeros(cout, device=d)
    ops.conv3x3_thin(x, w, b, stride, 0.1, reps=6)
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    nb = (N * H * W * cin + N * ho * wo * cout) * 2
    t = ops.conv3x3_thin.last_ms * 1e-3
    print(f"{name:10s} {cin}->{cout} s{stride}: {t*1e6:7.1f} us/launch (events, back to back) {nb/t/1e9:7.0f} GB/s")
