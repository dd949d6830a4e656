"""video / pair bf16 plans with and without the stride-2 WM path: where do they differ?  python tools/dbg_s2_plan.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200.models.PWCNet import PWCDCNet
dev = torch.device("cuda", 0)
torch.manual_seed(3)
B, H, W = 3, 128, 192
mods = {}
for s2 in ("1", "0"):
    os.environ["PWC_TC_WM_S2"] = s2
    torch.manual_seed(3)
    m = PWCDCNet(conv_mode="bf16").eval().to(dev)
    for video in (True, False):
        plan = m._plan(B, H, W, dev, video=video)
        plan.pack_weights(m)
        mods[(s2, video)] = (m, plan)
for video in (True, False):
    pa, pb = mods[("1", video)][1], mods[("0", video)][1]
    for seed in range(8):
        g = torch.Generator(device="cpu").manual_seed(seed)
        x = torch.rand(*pa.x_static.shape, generator=g).to(dev)
        outs = []
        for rep in range(3):
            oa = pa.run(x).clone(); ob = pb.run(x).clone()
            torch.cuda.synchronize()
            outs.append((oa, ob))
        d = [(oa - ob).abs().max().item() for oa, ob in outs]
        lv = {l: (pa.pa[l].float() - pb.pa[l].float()).abs().max().item() for l in range(1, 7)}
        print("video" if video else "pair ", "seed", seed, "flow diff per rep", d, "pyramid diff", {k: round(v, 4) for k, v in lv.items()})
