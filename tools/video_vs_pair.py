import os, sys, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
warnings.filterwarnings("ignore")
import torch
from oracle import pwcnet_oracle as O
from opticalflow_b200.models.PWCNet import PWCDCNet
mode = sys.argv[1] if len(sys.argv) > 1 else "mixed"
B, H, W = 3, int(sys.argv[2]) if len(sys.argv) > 2 else 256, int(sys.argv[3]) if len(sys.argv) > 3 else 448
torch.manual_seed(0)
m = PWCDCNet(conv_mode=mode).eval().cuda()
sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
g = torch.Generator().manual_seed(11)
frames = torch.rand(B + 1, 3, H, W, generator=g)
pairs = torch.cat([frames[:-1], frames[1:]], 1).contiguous()
ref = O.forward(sd, pairs)
a = m(pairs.cuda()).cpu()
plan = m._plan(B, H, W, torch.device("cuda"), video=True)
plan.pack_weights(m)
b = plan.run(frames.cuda()).clone().cpu()
epe = lambda u, v: 20 * (u - v).pow(2).sum(1).sqrt().mean(dim=(1, 2))
print(mode, "pair plan vs oracle", epe(a, ref).tolist())
print(mode, "video plan vs oracle", epe(b, ref).tolist())
print(mode, "video vs pair", epe(a, b).tolist())
