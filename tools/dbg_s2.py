"""stride-2 WM kernel (kMode 5) at the small pipeline shapes, repeated: python tools/dbg_s2.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from opticalflow_b200 import ops
d = torch.device("cuda")
for (cin, cout, H, W) in [(32, 64, 32, 48), (64, 96, 16, 24), (96, 128, 8, 12), (16, 32, 64, 96)]:
    for N in (4, 6):
        worst = 0.0
        for seed in range(6):
            torch.manual_seed(seed)
            x = torch.randn(N, H, W, cin, device=d).bfloat16()
            w = (torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)).bfloat16().float()
            b = torch.randn(cout, device=d) * 0.1
            ref = F.leaky_relu(F.conv2d(x.float().permute(0, 3, 1, 2), w, b, stride=2, padding=1), 0.1)
            outs = []
            for rep in range(3):
                o = ops.conv3x3_tc(x, w, b, 2, 1, 0.1)
                outs.append(o.float().permute(0, 3, 1, 2)[:, :cout])
            e = max((o - ref).abs().max().item() for o in outs) / ref.abs().max().item()
            same = all(torch.equal(outs[0], o) for o in outs[1:])
            worst = max(worst, e)
            if e > 1e-2 or not same:
                bad = (outs[0] - ref).abs().amax(dim=(1,))  # [N,Ho,Wo]
                idx = (bad > 1e-2 * ref.abs().max()).nonzero()
                print("  BAD", cin, cout, H, W, "N", N, "seed", seed, "rel err", e, "same", same, "first bad", idx[:6].tolist(), "count", len(idx))
        print(cin, cout, H, W, "N", N, "worst rel err", worst)
