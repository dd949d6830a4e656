"""GPU check of the tcgen05 conv against torch (bf16-rounded operands, fp32 math).  Each case runs in its own
subprocess under a timeout so that a hang cannot take the box down.   python tools/check_tc_conv.py [case_idx]"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CASES = [
    # cin, cout, stride, dil, N, H, W, out_f32
    (64, 128, 1, 1, 1, 8, 16, 0), (64, 128, 1, 1, 2, 24, 48, 0), (128, 128, 1, 1, 2, 21, 37, 0),
    (32, 32, 1, 1, 2, 16, 32, 0), (16, 16, 1, 1, 2, 16, 32, 0), (96, 64, 1, 16, 1, 40, 64, 0),
    (128, 96, 1, 8, 1, 24, 40, 0), (576, 128, 1, 1, 1, 16, 32, 0), (208, 208, 1, 1, 2, 7, 16, 0),
    (576, 2, 1, 1, 1, 16, 32, 1), (32, 2, 1, 1, 2, 16, 32, 1), (480, 64, 1, 1, 1, 16, 16, 0),
    (64, 96, 2, 1, 2, 32, 64, 0), (16, 32, 2, 1, 1, 32, 32, 0),
]


def run_case(i):
    import torch, torch.nn.functional as F
    from opticalflow_b200 import ops
    cin, cout, stride, dil, N, H, W, of32 = CASES[i]
    g = torch.Generator().manual_seed(100 + i)
    x = torch.randn(N, cin, H, W, generator=g).bfloat16()
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).bfloat16().float()
    b = torch.randn(cout, generator=g) * 0.1
    slope = 1.0 if of32 else 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w, b, stride=stride, padding=dil, dilation=dil), slope)
    d = torch.device("cuda")
    out = ops.conv3x3_tc(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), stride, dil, slope, out_f32=bool(of32))
    out = out.float().permute(0, 3, 1, 2).cpu()[:, :cout]
    err = (out - ref).abs().max().item(); sc = ref.abs().max().item()
    tol = (2e-3 if of32 else 1e-2) * sc
    print(f"case {i} {CASES[i]}: max err {err:.3e} (ref max {sc:.3f}) {'OK' if err <= tol else 'FAIL'}", flush=True)
    return err <= tol


if __name__ == "__main__":
    if len(sys.argv) > 1:
        sys.exit(0 if run_case(int(sys.argv[1])) else 1)
    bad = 0
    for i in range(len(CASES)):
        try:
            r = subprocess.run([sys.executable, __file__, str(i)], timeout=90, capture_output=True, text=True)
            sys.stdout.write(r.stdout[-400:])
            if r.returncode != 0:
                bad += 1
                if not r.stdout.strip():
                    print(f"case {i} {CASES[i]}: crashed rc={r.returncode}: {r.stderr[-300:]}")
        except subprocess.TimeoutExpired:
            bad += 1
            print(f"case {i} {CASES[i]}: TIMEOUT (hang)")
            break   # a hung kernel may leave the GPU busy: stop here
    print(f"{bad} failing")
