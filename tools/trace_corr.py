"""In-kernel timeline of the tensor-core cost-volume kernel (CTA 0): python tools/trace_corr.py [C H W B]
Also prints CUDA-event timings of the kernel alone (bf16 NHWC, concat-slice pitches)."""
import os, sys, ctypes
os.environ["PWC_CORR_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops, _lib
C, H, W, B = ([int(a) for a in sys.argv[1:5]] + [32, 112, 256, 8][len(sys.argv) - 1:])[:4]
d = torch.device("cuda")
torch.manual_seed(0)
cat = torch.randn(B, H, W, 544 + C, device=d).bfloat16()
f2 = torch.randn(B, H, W, C, device=d).bfloat16()
flow = 2 * torch.randn(B, H, W, 2, device=d)
f1, out = cat[..., 544:544 + C], cat[..., 448:529]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=d)
ts = []
for i in range(6):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.warp_corr_lrelu(f1, f2, flow, 5.0, layout="nhwc", out=out); e1.record(); e1.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
alg = B * H * W * (2 * C * 2 + 81 * 2 + 8)
print(f"C={C} {H}x{W} B={B}: us {['%.1f' % t for t in ts]}  -> {alg / min(ts[1:]) / 1e3:.0f} GB/s algorithmic")
n = 3 * 64 * 4
buf = (ctypes.c_ulonglong * n)()
_lib.check(_lib.load().pwc_debug_corr_trace(buf, n), "trace")
t0 = min(v for v in buf if v)
names = {0: ("producer", ["start", "table", "slot_free", "done"]), 1: ("mma", ["start", "a_full", "dy0", "dy8"]),
         2: ("epilogue", ["start", "dy0_full", "dy8_full", "written"])}
for role in range(3):
    nm, marks = names[role]
    print(f"--- {nm}: " + " ".join(f"{m:>10s}" for m in marks))
    for i in range(48):
        v = [buf[(role * 64 + i) * 4 + k] for k in range(4)]
        if not any(v): continue
        print(f"  {i:3d}     " + " ".join(f"{(x - t0) if x else -1:10d}" for x in v))
