"""Experiment: one batch-8 plan vs two batch-4 plans captured on two streams of ONE CUDA graph (latency-bound small-level
kernels of one half overlap the other half's)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200.models.PWCNet import PWCDCNet
from opticalflow_b200.engine_bf16 import PlanBF16
dev = torch.device("cuda")
torch.manual_seed(0)
m = PWCDCNet(conv_mode="bf16").eval().to(dev)
H, W = 448, 1024
x = torch.rand(8, 6, H, W, device=dev)

def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / n

p8 = PlanBF16(m, 8, H, W, dev); p8.pack_weights(m)
ref = p8.run(x).clone()
t8 = timeit(lambda: p8.run())
print(f"single plan B=8: {t8:.3f} ms -> {8 / t8 * 1e3:.0f} pairs/s")
for nsplit in (2, 4):
    bs = 8 // nsplit
    plans = [PlanBF16(m, bs, H, W, dev, use_graph=False) for _ in range(nsplit)]
    for i, p in enumerate(plans):
        p.pack_weights(m); p.x_static.copy_(x[i * bs:(i + 1) * bs]); p._launch_all()
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in range(nsplit)]
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        cur = torch.cuda.current_stream()
        for s, p in zip(streams, plans):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                p._launch_all()
        for s in streams:
            cur.wait_stream(s)
    t = timeit(lambda: g.replay())
    out = torch.cat([p.out for p in plans])
    print(f"{nsplit} plans of B={bs} on {nsplit} streams: {t:.3f} ms -> {8 / t * 1e3:.0f} pairs/s; max diff vs single {float((out - ref).abs().max()):.3g}")
