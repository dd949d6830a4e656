"""Finer CPU error budget (see tools/error_budget.py): mantissa-bit quantiser per tensor class."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import pwcnet_oracle as O
from tools.error_budget import run, epe, level_of, sd

H, W = 448, 1024
hf = torch.float16
inputs = {"rand": O.synthetic_input(1, H, W, 1), "shift": O.shifted_pair_input(1, H, W, 2)}
def is_pyr(n): return n.startswith("conv") and "_" not in n
def pyr_level(n): return int(n[4])
def is_dec(n): return ("_" in n and n.startswith("conv"))
base = lambda n, k: hf
def mk(f):
    return lambda n, k: (None if f(n, k) else hf)
policies = {
  "fp16 all": base,
  "A: pyramid fp32": mk(lambda n,k: is_pyr(n)),
  "A + L6 dec fp32": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)==6) or n=="lvl6"),
  "A + L6,5 dec fp32": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)>=5) or n in("lvl6","lvl5")),
  "A + L6-4 dec fp32": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)>=4) or n in("lvl6","lvl5","lvl4")),
  "A + L6-4 dec fp32 (corr/upfeat still fp16)": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)>=4)),
  "A + L6-4 dec weights fp32 only": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)>=4 and k=='w')),
  "A + L6-4 dec acts fp32 only": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)>=4 and k!='w') or n in("lvl6","lvl5","lvl4")),
  "A + L6-3 dec fp32": mk(lambda n,k: is_pyr(n) or (is_dec(n) and level_of(n)>=3) or n in("lvl6","lvl5","lvl4","lvl3")),
  "B: all dec+ctx fp32, pyramid fp16": mk(lambda n,k: not is_pyr(n)),
  "B + pyr weights fp32": mk(lambda n,k: (not is_pyr(n)) or k=='w'),
  "B + pyr acts fp32": mk(lambda n,k: (not is_pyr(n)) or k!='w'),
  "B + pyr L1 fp32": mk(lambda n,k: (not is_pyr(n)) or pyr_level(n)<=1),
  "B + pyr L1-2 fp32": mk(lambda n,k: (not is_pyr(n)) or pyr_level(n)<=2),
  "B + pyr L3-6 fp32": mk(lambda n,k: (not is_pyr(n)) or pyr_level(n)>=3),
  "B + pyr L4-6 fp32": mk(lambda n,k: (not is_pyr(n)) or pyr_level(n)>=4),
  "B + pyr a/aa fp32 (b layers fp16 out)": mk(lambda n,k: (not is_pyr(n)) or not (n.endswith("b") and k=='act')),
}
if len(sys.argv) > 1:
    policies = {k: v for k, v in policies.items() if any(s in k for s in sys.argv[1:])}
for iname, x in inputs.items():
    ref = O.forward(sd, x)
    print(f"[{iname}]")
    for pname, pol in policies.items():
        e = epe(run(x, pol), ref)
        print(f"  {pname:48s} = {20*e:.4f} px", flush=True)
