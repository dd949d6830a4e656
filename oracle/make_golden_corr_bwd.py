"""Generates tests/golden/corr_bwd.npz: gradients of the UNMODIFIED reference correlation (its pure-torch
shift-and-dot, correlation.py:12-42, divided by C as the CUDA kernel does -- oracle/ref_harness.py) by torch autograd.
The reference's own backward is CUDA only (correlation_cuda_kernel.cu:150-334) and cannot run in the build container;
autograd of its forward formula is the same linear map.
    python oracle/make_golden_corr_bwd.py        (needs /root/reference; run in the build container only)
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pwcnet_oracle as O  # noqa: E402
from oracle import ref_harness  # noqa: E402

CASES = [(3, 2, 14, 32, 24), (6, 1, 7, 16, 40), (0, 2, 5, 6, 3)]      # (level seed, B, H, W, C); the last is smaller than 9x9


def inputs(lvl, B, H, W, C):
    f1, f2, _ = O.corr_microbench_inputs(lvl, B, H, W, C)
    g = torch.Generator().manual_seed(40 + lvl)
    return f1, f2, torch.randn(B, 81, H, W, generator=g)


def main():
    _, corr_mod = ref_harness.load_reference()
    out = {"cases": np.array(CASES, dtype=np.int64)}
    for i, (lvl, B, H, W, C) in enumerate(CASES):
        f1, f2, go = inputs(lvl, B, H, W, C)
        a, b = f1.double().requires_grad_(True), f2.double().requires_grad_(True)
        y = corr_mod.correlation_onnx(a, b, 4, 1, 4, 1, 1, 1)
        y.backward(go.double())
        out[f"g1_{i}"] = a.grad.float().numpy()
        out[f"g2_{i}"] = b.grad.float().numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "corr_bwd.npz"), **out)
    print("wrote tests/golden/corr_bwd.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
