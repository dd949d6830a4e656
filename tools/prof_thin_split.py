"""A few launches of the split thin conv, for ncu: python tools/prof_thin_split.py cin cout N H W [stride]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, N, H, W = [int(v) for v in sys.argv[1:6]]
stride = int(sys.argv[6]) if len(sys.argv) > 6 else 1
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d)
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
with ops.half_format("f16"):
    ops.conv3x3_thin_split(x, w, b, reps=4, stride=stride)
print(f"{ops.conv3x3_thin_split.last_ms * 1e3:.1f} us per launch")
