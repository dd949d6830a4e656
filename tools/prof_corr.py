"""Run the fused warp+corr kernel a few times at one pyramid level (for ncu): python tools/prof_corr.py [level] [B] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
from oracle import pwcnet_oracle as O
lvl = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
C, H, W = {6: (196, 7, 16), 5: (128, 14, 32), 4: (96, 28, 64), 3: (64, 56, 128), 2: (32, 112, 256)}[lvl]
d = torch.device("cuda")
f1, f2, flow = O.corr_microbench_inputs(lvl, B, H, W, C, with_flow=(lvl < 6))
nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(d)
a, b = nhwc(f1), nhwc(f2)
fl = None if flow is None else nhwc(flow)
out = torch.empty(B, H, W, 81, device=d)
sc = O.FLOW_SCALE.get(lvl, 1.0)
ops.warp_corr_lrelu(a, b, fl, sc, layout="nhwc", out=out)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(reps):
        ops.warp_corr_lrelu(a, b, fl, sc, layout="nhwc", out=out)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g.replay(); torch.cuda.synchronize()
e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) * 1e-3 / reps
nb = 4.0 * B * H * W * (2 * C + 81 + (2 if lvl < 6 else 0))
print(f"L{lvl} B{B}: {t*1e6:.1f} us/launch (graph, L2-warm) -> {nb/t/1e9:.0f} GB/s algorithmic")
