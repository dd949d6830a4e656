"""Per-launch CUDA-event times of the bf16 plan: python tools/layer_times.py [B] [H] [W] [reps]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import _lib
from opticalflow_b200.models.PWCNet import PWCDCNet
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
H = int(sys.argv[2]) if len(sys.argv) > 2 else 448
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
torch.manual_seed(0)
m = PWCDCNet(conv_mode=os.environ.get("PWC_MODE", "bf16")).eval().cuda()
dev = torch.device("cuda")
plan = m._plan(B, H, W, dev)
plan.pack_weights(m)
plan.x_static.copy_(torch.rand(B, 6, H, W, device=dev))
plan._launch_all(); torch.cuda.synchronize()
rows = plan.time_ops(reps=reps)          # each launch captured 8x into its own CUDA graph: GPU time without host launch cost
tot = [t for _, t, _ in rows]
T = sum(tot)
print(f"B={B} {W}x{H}: sum of launches {T*1e3:.3f} ms -> {B/T:.1f} pairs/s (each timed alone, back to back with itself)")
for name, t, fl in rows:
    print(f"{name:18s} {t*1e6:9.1f} us  {100*t/T:5.1f}%  {fl/t/1e12 if t > 0 else 0:8.1f} TFLOP/s")
# whole-graph time
for _ in range(3): plan.run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): plan.run()
e1.record(); e1.synchronize()
tg = e0.elapsed_time(e1) * 1e-4
print(f"graph replay: {tg*1e3:.3f} ms/step -> {B/tg:.1f} pairs/s")
