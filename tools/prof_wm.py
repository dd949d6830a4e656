"""A few launches of one WM tcgen05 conv, for ncu: python tools/prof_wm.py cin cout dil N H W"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, dil, N, H, W = [int(v) for v in sys.argv[1:7]]
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
ops.conv3x3_tc(x, w, b, 1, dil, 0.1, reps=3)
print(f"{ops.conv3x3_tc.last_ms * 1e3:.1f} us per launch")
