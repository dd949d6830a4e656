"""One bf16 forward (batch 8, 1024x448) inside a cudaProfilerStart/Stop range, launch by launch (no graph), for
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file X python tools/prof_forward.py
Writes the launch order (layer names) to gpurun_out/forward_launch_names.txt so the CSV rows can be labelled."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import _lib
from opticalflow_b200.models.PWCNet import PWCDCNet
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
H = int(sys.argv[2]) if len(sys.argv) > 2 else 448
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
torch.manual_seed(0)
m = PWCDCNet(conv_mode=os.environ.get("PWC_MODE", "bf16")).eval().cuda()
dev = torch.device("cuda")
plan = m._plan(B, H, W, dev)
plan.pack_weights(m)
plan.x_static.copy_(torch.rand(B, 6, H, W, device=dev))
plan._launch_all(); plan._launch_all(); torch.cuda.synchronize()
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/forward_launch_names.txt", "w") as f:
    for name, fl in plan.op_info:
        f.write(f"{name} {fl:.0f}\n")
torch.cuda.cudart().cudaProfilerStart()
plan._launch_all()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok", len(plan.ops), "launches")
