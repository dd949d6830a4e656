"""CPU error budget of reduced-precision storage (oracle with a per-layer quantiser): which roundings produce the EPE delta
of the 16-bit conv modes.  Not on the product path.  python tools/error_budget.py [H W]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import pwcnet_oracle as O

H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (448, 1024)
torch.set_num_threads(8)
sd = O.reference_like_state_dict(0)


def level_of(name):
    if name.startswith("dc_conv"): return 2
    if name.startswith("predict_flow") or name.startswith("upfeat") or name.startswith("deconv"): return int(name[-1])
    return int(name[4])


def run(x, policy):
    """policy(name, kind) -> dtype or None; kind in {'w','act','corr','upflow','upfeat'}"""
    def q(t, name, kind):
        dt = policy(name, kind)
        return t if dt is None else t.to(dt).float()
    orig_conv, orig_q = O._conv, O._q
    def conv(sd_, name, x_, stride=1, pad=1, dil=1, act=True):
        key = f"{name}.0" if act else name
        w = sd_[f"{key}.weight"]
        if name != "conv1a":
            w = q(w, name, 'w')
        y = F.conv2d(x_, w, sd_[f"{key}.bias"], stride=stride, padding=pad, dilation=dil)
        return q(F.leaky_relu(y, 0.1), name, 'act') if act else y
    cur = [6]
    def qq(t):
        return q(t, f"lvl{cur[0]}", 'misc')
    O._conv = conv
    O._q = qq
    O._BF16_SIM[0] = True
    try:
        return O._forward(sd, x, False, False)
    finally:
        O._conv, O._q = orig_conv, orig_q
        O._BF16_SIM[0] = False


def epe(a, b):
    return (a - b).pow(2).sum(1).sqrt().mean().item()


inputs = {"rand": O.synthetic_input(1, H, W, 1), "shift": O.shifted_pair_input(1, H, W, 2)}
bf, hf = torch.bfloat16, torch.float16
policies = {
    "all bf16": lambda n, k: bf,
    "all fp16": lambda n, k: hf,
    "fp16, weights fp32": lambda n, k: None if k == 'w' else hf,
    "fp16 acts fp32, weights fp16": lambda n, k: hf if k == 'w' else None,
    "bf16 pyramid only": lambda n, k: bf if (n.startswith("conv") and "_" not in n) else None,
    "fp16 pyramid only": lambda n, k: hf if (n.startswith("conv") and "_" not in n) else None,
    "fp16 decoders only": lambda n, k: hf if ("_" in n or n.startswith("lvl")) else None,
    "fp16 L2+L3 (dec+ctx), fp32 pyramid & L4-6": lambda n, k: hf if (("_" in n) and level_of(n) <= 3) else None,
    "fp16 all but L4-6 decoders": lambda n, k: None if (("_" in n and not n.startswith("dc") and level_of(n) >= 4) or n in ("lvl6","lvl5","lvl4")) else hf,
}
if __name__ == "__main__":
  for iname, x in inputs.items():
      t0 = time.time()
      ref = O.forward(sd, x)
      print(f"[{iname}] ref forward {time.time()-t0:.1f}s |flow2| max {ref.abs().max():.3f} (x20 = {20*ref.abs().max():.1f} px)")
      for pname, pol in policies.items():
          out = run(x, pol)
          e = epe(out, ref)
          print(f"  {pname:48s} EPE delta {e:.5f} (px/20)  = {20*e:.4f} px")
