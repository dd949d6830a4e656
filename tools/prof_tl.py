"""A few launches of the tap-lane split conv, for ncu: python tools/prof_tl.py C N H W"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
C, N, H, W = [int(v) for v in sys.argv[1:5]]
d = torch.device("cuda")
x = torch.randn(N, H, W, C, device=d)
w = torch.randn(C, C, 3, 3, device=d) / (3 * C ** 0.5)
b = torch.zeros(C, device=d)
with ops.half_format("f16"):
    ops.conv3x3_tc_split(x, w, b, tl=True, reps=3)
print(f"{ops.conv3x3_tc_split.last_ms * 1e3:.1f} us per launch")
