"""A few launches of the stride-2 split conv, for ncu: python tools/prof_s2.py cin cout N H W"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, N, H, W = [int(v) for v in sys.argv[1:6]]
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d)
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
with ops.half_format("f16"):
    ops.conv3x3_tc_split(x, w, b, stride=2, reps=3)
print(f"{ops.conv3x3_tc_split.last_ms * 1e3:.1f} us per launch")
