"""Run one tensor-core conv shape a few times (for ncu): python tools/prof_conv.py cin cout dil N H W [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, dil, N, H, W = [int(a) for a in sys.argv[1:7]]
reps = int(sys.argv[7]) if len(sys.argv) > 7 else 4
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
ops.conv3x3_tc(x, w, b, 1, dil, 0.1, reps=reps)
ms = ops.conv3x3_tc.last_ms
print(f"{cin}->{cout} d{dil} N{N} {W}x{H}: {ms*1e3:.1f} us/launch, {2.0*N*H*W*cout*cin*9/ms/1e9:.1f} TFLOP/s")
