"""GPU check of the halo tcgen05 conv against torch, over env-selected variants.  Each (variant, case) runs in its own
subprocess under a timeout so that a hang cannot take the box down.  python tools/check_tc_halo.py"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CASES = [
    # cin, cout, stride, dil, N, H, W, out_f32
    (64, 128, 1, 1, 1, 16, 16, 0), (64, 128, 1, 1, 2, 40, 72, 0), (128, 128, 1, 2, 2, 21, 37, 0),
    (576, 128, 1, 1, 1, 32, 48, 0), (544, 32, 1, 1, 1, 28, 64, 0), (128, 96, 1, 8, 1, 40, 64, 0),
    (128, 128, 1, 4, 1, 56, 64, 0), (576, 2, 1, 1, 1, 16, 32, 1), (208, 208, 1, 1, 2, 7, 16, 0),
    (480, 64, 1, 1, 8, 112, 256, 0),
    (32, 32, 1, 1, 2, 24, 40, 0), (16, 16, 1, 1, 2, 33, 48, 0),
]
VARIANTS = [
    {}, {"PWC_TC_WM_ROWS": "8", "PWC_TC_HALO_RESIDENT": "0"}, {"PWC_TC_WM_ROWS": "16"}, {"PWC_TC_WM_ROWS": "32"},
    {"PWC_TC_WM": "0"}, {"PWC_TC_WM": "0", "PWC_TC_HALO_HALVES": "1", "PWC_TC_HALO_NSPLIT": "2"},
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
    hung = False
    for v in VARIANTS:
        print("=== variant", v, flush=True)
        env = dict(os.environ); env.update(v)
        bad = 0
        for i in range(len(CASES)):
            try:
                r = subprocess.run([sys.executable, __file__, str(i)], timeout=90, capture_output=True, text=True, env=env)
                sys.stdout.write(r.stdout[-300:])
                if r.returncode != 0:
                    bad += 1
                    if not r.stdout.strip():
                        print(f"case {i} {CASES[i]}: crashed rc={r.returncode}: {r.stderr[-300:]}")
            except subprocess.TimeoutExpired:
                bad += 1; hung = True
                print(f"case {i} {CASES[i]}: TIMEOUT (hang)")
                break
        print(f"--- {bad} failing", flush=True)
        if hung:
            break
