"""In-kernel timeline of the halo conv (CTA 0): python tools/trace_conv.py cin cout dil N H W"""
import os, sys, ctypes
os.environ["PWC_TC_HALO_BO"] = str(16 | int(os.environ.get("PWC_TC_HALO_BO", "0")))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops, _lib
cin, cout, dil, N, H, W = [int(a) for a in sys.argv[1:7]]
d = torch.device("cuda")
x = torch.randn(N, H, W, cin, device=d).bfloat16()
w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
b = torch.zeros(cout, device=d)
ops.conv3x3_tc(x, w, b, 1, dil, 0.1, reps=3)
buf = (ctypes.c_ulonglong * 8192)()
_lib.check(_lib.load().pwc_debug_halo_trace(buf, 8192), "trace")
def dump(base, name, limit=140):
    ev = [(buf[base + i] >> 56, buf[base + i] & ((1 << 56) - 1)) for i in range(4096) if buf[base + i]]
    if not ev: return
    t0 = ev[0][1]; prev = t0
    print(f"--- {name}: {len(ev)} events")
    for tag, t in ev[:limit]:
        print(f"  tag {tag:2d}  t={t - t0:8d}  dt={t - prev:6d}")
        prev = t
dump(0, "MMA thread (1 item start, 2 A ready, 3 B ready, 4 tap issued)")
dump(4096, "producer (10 A slot free, 12 B slot free)", 60)
