"""Run the fused warp+corr kernel a few times at one pyramid level (for ncu / timing):
    python tools/prof_corr.py [level] [B] [reps] [f32|bf16]
bf16: NHWC bf16 features, bf16 output inside an 88-channel pitch (as in the concat buffer) -> tensor-core / small kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import ops
from oracle import pwcnet_oracle as O
lvl = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
bf = len(sys.argv) > 4 and sys.argv[4] == "bf16"
C, H, W = {6: (196, 7, 16), 5: (128, 14, 32), 4: (96, 28, 64), 3: (64, 56, 128), 2: (32, 112, 256)}[lvl]
d = torch.device("cuda")
f1, f2, flow = O.corr_microbench_inputs(lvl, B, H, W, C, with_flow=(lvl < 6))
nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(d)
a, b = nhwc(f1), nhwc(f2)
if bf:
    aa = torch.zeros(B, H, W, (C + 15) // 16 * 16, device=d, dtype=torch.bfloat16)
    aa[..., :C] = a
    a = aa[..., :C]
    Cp = (C + 15) // 16 * 16                     # image-2 features with the pyramid buffers' padded pitch
    bb = torch.zeros(B, H, W, Cp, device=d, dtype=torch.bfloat16)
    bb[..., :C] = b
    b = bb[..., :C]
fl = None if flow is None else nhwc(flow)
out = torch.empty(B, H, W, 88, device=d, dtype=a.dtype)[..., :81] if bf else torch.empty(B, H, W, 81, device=d)
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
e = 2.0 if bf else 4.0
nb = B * H * W * (2 * C * e + 81 * e + (8 if lvl < 6 else 0))
print(f"L{lvl} B{B} {'bf16' if bf else 'f32'}: {t*1e6:.1f} us/launch (graph, back to back) -> {nb/t/1e9:.0f} GB/s algorithmic")
