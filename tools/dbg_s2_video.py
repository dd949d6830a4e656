"""VideoFlowPipeline with / without the stride-2 WM path on the frames of tests/test_gpu_driver.py: python tools/dbg_s2_video.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import driver
from opticalflow_b200.models.PWCNet import PWCDCNet
dev = torch.device("cuda", 0)
B, H, W, N = 3, 100, 150, 8
g = torch.Generator().manual_seed(9)
frames = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, generator=g).pin_memory()
res = {}
for full_res in (False, True):
    for s2 in ("1", "0"):
        os.environ["PWC_TC_WM_S2"] = s2
        torch.manual_seed(3)
        m = PWCDCNet(conv_mode="bf16").eval().to(dev)
        pipe = driver.VideoFlowPipeline(m, B, H, W, dev, swap_rb=True, mult=20.0, full_res=full_res, share_pyramid=True)
        for rep in range(2):
            got = {}
            pipe.run(frames, sink=lambda t0, c, f: got.update({t0 + k: f[k].clone() for k in range(c)}))
            res[(full_res, s2, rep)] = got
    for t in range(N - 1):
        a, b = res[(full_res, "1", 0)][t], res[(full_res, "0", 0)][t]
        a2 = res[(full_res, "1", 1)][t]
        d = (a - b)
        print("full_res", full_res, "pair", t, "S2 on vs off: max", round(d.abs().max().item(), 4), "mean EPE", round(d.pow(2).sum(0).sqrt().mean().item(), 4),
              "| run-to-run max", (a - a2).abs().max().item(), "| flow max", round(b.abs().max().item(), 1))
