"""EPE delta of a conv mode vs the CPU oracle at 1024x448 (both synthetic inputs): python tools/mixed_epe.py [mode]"""
import os, sys, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
warnings.filterwarnings("ignore")
import torch
from oracle import pwcnet_oracle as O
from opticalflow_b200.models.PWCNet import PWCDCNet
mode = sys.argv[1] if len(sys.argv) > 1 else "mixed"
torch.manual_seed(0)
m = PWCDCNet(conv_mode=mode).eval().cuda()
sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
for name, x in (("rand", O.synthetic_input(1, 448, 1024, 1)), ("shift", O.shifted_pair_input(1, 448, 1024, 3)), ("rand2", O.synthetic_input(1, 448, 1024, 5))):
    ref = O.forward(sd, x)
    out = m(x.cuda()).cpu()
    e = 20 * (out - ref).pow(2).sum(1).sqrt().mean().item()
    print(f"{mode} levels={os.environ.get('PWC_MIXED_LEVELS','6,5,4')} {name}: mean EPE delta {e:.4f} px")
