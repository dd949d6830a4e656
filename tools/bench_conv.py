"""Time one tensor-core conv shape under env-selected kernel variants (each in a subprocess).
python tools/bench_conv.py cin cout dil N H W"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
VARIANTS = [{"PWC_TC_HALO": "0"}, {"PWC_TC_WM": "0"}, {}, {"PWC_TC_WM_ROWS": "16"}]
if __name__ == "__main__":
    if os.environ.get("BENCH_CONV_CHILD"):
        import torch
        from opticalflow_b200 import ops
        cin, cout, dil, N, H, W = [int(a) for a in sys.argv[1:7]]
        d = torch.device("cuda")
        x = torch.randn(N, H, W, cin, device=d).bfloat16()
        w = torch.randn(cout, cin, 3, 3, device=d) / (3 * cin ** 0.5)
        b = torch.zeros(cout, device=d)
        ops.conv3x3_tc(x, w, b, 1, dil, 0.1, reps=11)
        ms = ops.conv3x3_tc.last_ms
        fl = 2.0 * N * H * W * cout * cin * 9
        print(f"{ms*1e3:9.1f} us  {fl/ms/1e9:8.1f} TFLOP/s")
        sys.exit(0)
    for v in VARIANTS:
        env = dict(os.environ); env.update(v); env["BENCH_CONV_CHILD"] = "1"
        try:
            r = subprocess.run([sys.executable, __file__] + sys.argv[1:7], timeout=120, capture_output=True, text=True, env=env)
            print(f"{' '.join(sys.argv[1:7]):28s} {str(v):46s} {r.stdout.strip()[-40:]} {r.stderr.strip()[-200:] if r.returncode else ''}", flush=True)
        except subprocess.TimeoutExpired:
            print(v, "TIMEOUT"); break
