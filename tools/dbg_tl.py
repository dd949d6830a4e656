import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from opticalflow_b200 import ops
cin, cout, N, H, W = [int(v) for v in sys.argv[1:6]] if len(sys.argv) > 5 else (128, 32, 1, 24, 16)
g = torch.Generator().manual_seed(1)
x = torch.randn(N, cin, H, W, generator=g).bfloat16()
w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).bfloat16().float()
b = torch.randn(cout, generator=g) * 0.1
ref = F.leaky_relu(F.conv2d(x.float(), w, b, padding=1), 0.1)
d = torch.device("cuda:0")
out = ops.conv3x3_tc(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), 1, 1, 0.1, tl=True)
got = out.float().permute(0, 3, 1, 2).cpu()[:, :cout]
print("max err", (got - ref).abs().max().item(), "scale", ref.abs().max().item())
