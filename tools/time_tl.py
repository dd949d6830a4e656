"""time one tap-lane (or plain) WM conv: python tools/time_tl.py cin cout N H W [tl]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
cin, cout, N, H, W = [int(v) for v in sys.argv[1:6]]
tl = len(sys.argv) > 6 and sys.argv[6] == "1"
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
ops.conv3x3_tc(x, w, b, 1, 1, 0.1, reps=8, tl=tl)
print(f"{cin}->{cout} tl={tl}: {ops.conv3x3_tc.last_ms * 1e3:.1f} us per launch")
if os.environ.get("PWC_TC_HALO_BO") == "32":
    import ctypes
    from opticalflow_b200 import _lib
    n = 3 * 64 * 4
    buf = (ctypes.c_ulonglong * n)()
    _lib.check(_lib.load().pwc_debug_wm_trace(buf, n), "trace")
    t0 = min(v for v in buf if v)
    names = {0: ("producer", ["start", "x_empty", "-", "-"]), 1: ("mma", ["start", "t_empty", "x_full", "issued"]),
             2: ("epilogue", ["start", "t_full", "half0", "half1"])}
    for role in range(3):
        nm, marks = names[role]
        print(f"--- {nm}: " + " ".join(f"{m:>10s}" for m in marks))
        for i in range(8):
            v = [buf[(role * 64 + i) * 4 + k] for k in range(4)]
            if not any(v): continue
            print(f"  {i:3d}     " + " ".join(f"{(q - t0) if q else -1:10d}" for q in v))
