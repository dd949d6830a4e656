"""Timeline of CTA 0 of the stride-2 split conv in the WM kernel's pair view: python tools/trace_s2.py cin cout N H W"""
import os, sys, ctypes
os.environ["PWC_TC_HALO_BO"] = "32"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops, _lib
cin, cout, N, H, W = [int(v) for v in sys.argv[1:6]]
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d)
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
with ops.half_format("f16"):
    ops.conv3x3_tc_split(x, w, b, stride=2, reps=6)
print(f"{ops.conv3x3_tc_split.last_ms * 1e3:.1f} us per launch")
n = 3 * 64 * 4
buf = (ctypes.c_ulonglong * n)()
_lib.check(_lib.load().pwc_debug_wm_trace(buf, n), "trace")
vals = [v for v in buf if v]
if not vals:
    print("no trace (per-tap kernel)"); sys.exit(0)
t0 = min(vals)
names = {0: ("producer", ["start", "x_empty", "-", "-"]), 1: ("mma", ["start", "t_empty", "x_full", "issued"]),
         2: ("epilogue", ["start", "t_full", "half0", "half1"])}
for role in range(3):
    nm, marks = names[role]
    print(f"--- {nm}: " + " ".join(f"{m:>10s}" for m in marks))
    for i in range(14):
        v = [buf[(role * 64 + i) * 4 + k] for k in range(4)]
        if not any(v): continue
        print(f"  {i:3d}     " + " ".join(f"{(q - t0) if q else -1:10d}" for q in v))
