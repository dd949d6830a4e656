"""Timeline of CTA 0 of the WM tcgen05 conv: python tools/trace_wm.py cin cout dil N H W [s2d_g]
s2d_g > 1 runs the layer in the space-to-depth form the engine uses for thin pyramid layers."""
import os, sys, ctypes
os.environ["PWC_TC_HALO_BO"] = "32"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops, _lib
from opticalflow_b200.engine_bf16 import _s2d_weights
a = [int(v) for v in sys.argv[1:]]
cin, cout, dil, N, H, W = a[:6]
g = a[6] if len(a) > 6 else 1
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
if g > 1:
    w, b = _s2d_weights("k", g)({"k.weight": w, "k.bias": b})
    x = x.view(N, H, W // g, g * cin)
ops.conv3x3_tc(x, w, b, 1, dil, 0.1, reps=6)
print(f"{ops.conv3x3_tc.last_ms * 1e3:.1f} us per launch")
n = 3 * 64 * 4
buf = (ctypes.c_ulonglong * n)()
_lib.check(_lib.load().pwc_debug_wm_trace(buf, n), "trace")
t0 = min(v for v in buf if v)
names = {0: ("producer", ["start", "x_empty", "-", "-"]), 1: ("mma", ["start", "t_empty", "x_full", "issued"]),
         2: ("epilogue", ["start", "t_full", "half0", "half1"])}
for role in range(3):
    nm, marks = names[role]
    print(f"--- {nm}: " + " ".join(f"{m:>10s}" for m in marks))
    for i in range(30):
        v = [buf[(role * 64 + i) * 4 + k] for k in range(4)]
        if not any(v): continue
        print(f"  {i:3d}     " + " ".join(f"{(q - t0) if q else -1:10d}" for q in v))
