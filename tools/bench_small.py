"""Small-level layer shapes: tcgen05 conv vs the warp-MMA small-level conv, CUDA-event time per launch.
python tools/bench_small.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
d = torch.device("cuda")
cases = [  # name, cin, cout, stride, N, H, W
    ("conv6_0", 81, 128, 1, 8, 7, 16), ("conv6_1", 209, 128, 1, 8, 7, 16), ("conv6_2", 337, 96, 1, 8, 7, 16),
    ("conv6_3", 433, 64, 1, 8, 7, 16), ("conv6_4", 497, 32, 1, 8, 7, 16), ("conv6a/b", 196, 196, 1, 16, 7, 16),
    ("conv6aa", 128, 196, 2, 16, 14, 32), ("conv5aa/b", 128, 128, 1, 16, 14, 32), ("conv5a", 96, 128, 2, 16, 28, 64),
    ("conv5_0", 213, 128, 1, 8, 14, 32), ("conv5_1", 341, 128, 1, 8, 14, 32), ("conv5_2", 469, 96, 1, 8, 14, 32),
    ("conv5_3", 565, 64, 1, 8, 14, 32), ("conv5_4", 629, 32, 1, 8, 14, 32),
]
for name, cin, cout, stride, N, H, W in cases:
    span = (cin + 31) // 32 * 32
    x = torch.randn(N, H, W, span, device=d).bfloat16()
    w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
    b = torch.zeros(cout, device=d)
    ops.conv3x3_tc(x, w, b, stride, 1, 0.1, cin_span=span, reps=3)
    t_tc = ops.conv3x3_tc.last_ms * 1e3
    row = f"{name:10s} {cin:4d}->{cout:3d} s{stride} N{N} {W}x{H}: tc {t_tc:6.1f} us | mma"
    for cfg in ("-1", "0", "1", "2"):
        os.environ["PWC_MMA_SMALL_CFG"] = cfg
        # the env var is read once per process: pass through a fresh library load is not possible, so only -1 is live
        if cfg != "-1":
            break
        ops.conv3x3_mma_small(x, w, b, stride, 0.1, cin_span=span, reps=3)
        row += f" {ops.conv3x3_mma_small.last_ms * 1e3:6.1f} us"
    print(row)
