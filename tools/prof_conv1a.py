"""conv1a at the bench shape, a few launches (for ncu / timing): python tools/prof_conv1a.py [B] [H] [W]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from opticalflow_b200 import _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
H = int(sys.argv[2]) if len(sys.argv) > 2 else 448
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
d = torch.device("cuda")
x = torch.rand(B, 6, H, W, device=d)
w = torch.randn(16, 3, 3, 3, device=d) / 5
b = torch.zeros(16, device=d)
out = torch.empty(2 * B, H // 2, W // 2, 16, device=d, dtype=torch.bfloat16)
lib = _lib.load()
st = _lib.stream_handle()
for _ in range(3):
    lib.pwc_conv1a_bf16_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), B, H, W, 0.1, st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    lib.pwc_conv1a_bf16_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), B, H, W, 0.1, st)
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 10 * 1e-3
nb = x.numel() * 4 + out.numel() * 2
print(f"conv1a B{B} {W}x{H}: {t*1e6:.1f} us -> {nb/t/1e9:.0f} GB/s, {2.0*2*B*(H//2)*(W//2)*16*27/t/1e12:.1f} TFLOP/s")
