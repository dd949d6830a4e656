"""One thin-layer conv shape (for ncu): python tools/prof_thin.py cin cout stride N H W"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, stride, N, H, W = [int(a) for a in sys.argv[1:7]]
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
ops.conv3x3_thin(x, w, b, stride, 0.1, reps=3)
print("ok")
