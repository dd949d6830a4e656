"""Sustained pairs/s of the uint8 video path (BASELINE configs[4]: 1920x1080 stream, pwc_extract_flow_video.py) on ONE GPU:
pinned uint8 frames -> VideoFlowPipeline (H2D | pre-process + forward + post-process | D2H) -> flows on the host.
    python tools/bench_video.py [H] [W] [B] [n_frames] [full_res]
With torchrun each rank processes its own contiguous chunk of the stream (+1 overlapping frame), no collective."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import driver
from opticalflow_b200.models.PWCNet import PWCDCNet
H = int(sys.argv[1]) if len(sys.argv) > 1 else 1080
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
B = int(sys.argv[3]) if len(sys.argv) > 3 else 4
NF = int(sys.argv[4]) if len(sys.argv) > 4 else 65
full = len(sys.argv) > 5 and sys.argv[5] == "1"
rank, local, world = driver.dist_env()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
torch.manual_seed(0)
m = PWCDCNet(conv_mode="bf16").eval().to(dev)
g = torch.Generator().manual_seed(rank)
frames = torch.randint(0, 256, (NF, H, W, 3), dtype=torch.uint8, generator=g).pin_memory()
pipe = driver.VideoFlowPipeline(m, B, H, W, dev, full_res=full)
acc = []
pipe.run(frames[:2 * B + 1])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
n = pipe.run(frames, sink=lambda t0, c, f: acc.append(float(f[0, 0, 0, 0])))
e1.record(); e1.synchronize()
t = e0.elapsed_time(e1) * 1e-3
print(json.dumps({"workload": f"uint8 video {W}x{H} (padded to {pipe.Wp}x{pipe.Hp}), {NF} frames, batch {B} pairs, "
                              f"{'full-resolution' if full else 'quarter-resolution'} flow out", "rank": rank, "world": world,
                  "pairs": n, "pairs_per_s": n / t, "ms_per_pair": 1e3 * t / n,
                  "h2d_bytes_per_pair": H * W * 3 * (B + 1) / B, "d2h_bytes_per_pair": 8 * pipe.oh * pipe.ow}))
