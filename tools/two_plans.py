"""Throughput of two plans in flight on two streams against one (batch B each): python tools/two_plans.py [mode] [B]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import engine
from opticalflow_b200.models.PWCNet import PWCDCNet
mode = sys.argv[1] if len(sys.argv) > 1 else "mixed"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
H, W = 448, 1024
NPL = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = PWCDCNet(conv_mode=mode).eval().to(dev)
plans = []
for i in range(NPL):
    p = engine.make_plan(m, B, H, W, dev, mode, m.align_corners)
    p.pack_weights(m)
    plans.append(p)
x = torch.rand(B, 6, H, W, device=dev)
streams = [torch.cuda.Stream(dev) for _ in range(NPL)]
for p, s in zip(plans, streams):
    with torch.cuda.stream(s):
        for _ in range(3):
            p.run(x)
torch.cuda.synchronize()
ref = plans[0].run(x).clone()
assert torch.equal(ref, plans[1].run(x)), "the two plans disagree"
def timed(nplans, K=40):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams[:nplans]:
        s.wait_event(e0)
    for k in range(K):
        i = k % nplans
        with torch.cuda.stream(streams[i]):
            plans[i].graph.replay()
    for s in streams[:nplans]:
        torch.cuda.current_stream().wait_stream(s)
    e1.record(); e1.synchronize()
    return K * B / (e0.elapsed_time(e1) * 1e-3)
for n in list(range(1, NPL + 1)) * 2:
    print(f"{n} plan(s) in flight: {timed(n):.1f} pairs/s")
