"""One small-level conv shape on the warp-MMA kernel (for ncu): python tools/prof_small.py cin cout stride N H W"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, stride, N, H, W = [int(a) for a in sys.argv[1:7]]
d = torch.device("cuda")
span = (cin + 31) // 32 * 32
x = torch.randn(N, H, W, span, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
ops.conv3x3_mma_small(x, w, b, stride, 0.1, cin_span=span, reps=3)
print("ok")
