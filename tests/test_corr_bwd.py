"""Correlation backward (correlation_cuda_kernel.cu:150-334) and the standalone warp
(PWCNet.py:141-177): oracle vs goldens generated from the unmodified reference (oracle/make_golden_corr_bwd.py,
oracle/make_golden.py), CUDA kernels vs both."""
import os

import numpy as np
import pytest
import torch

from oracle import pwcnet_oracle as O

TOL = 1e-5          # relative to the tensor's max magnitude (fp32 sums of <= 81*C products, any order)


def _inputs(lvl, B, H, W, C):
    f1, f2, _ = O.corr_microbench_inputs(lvl, B, H, W, C)
    g = torch.Generator().manual_seed(40 + lvl)
    return f1, f2, torch.randn(B, 81, H, W, generator=g)


def _cases(golden_dir):
    g = np.load(os.path.join(golden_dir, "corr_bwd.npz"))
    return g, [tuple(int(v) for v in row) for row in g["cases"]]


def test_backward_oracle_matches_reference_autograd(golden_dir):
    g, cases = _cases(golden_dir)
    for i, (lvl, B, H, W, C) in enumerate(cases):
        f1, f2, go = _inputs(lvl, B, H, W, C)
        g1, g2 = O.correlation_backward_numpy(f1.numpy(), f2.numpy(), go.numpy())
        for mine, ref in ((g1, g[f"g1_{i}"]), (g2, g[f"g2_{i}"])):
            assert np.abs(mine - ref).max() <= 1e-6 * np.abs(ref).max()


def test_backward_oracle_is_adjoint_of_forward():
    """<corr(f1,f2), G> is bilinear: d/df1 <corr, G> . V == <corr(V, f2), G>."""
    f1, f2, go = _inputs(4, 1, 9, 11, 5)
    v = torch.randn(f1.shape, generator=torch.Generator().manual_seed(3))
    g1, g2 = O.correlation_backward_numpy(f1.numpy(), f2.numpy(), go.numpy())
    lhs1 = float((torch.from_numpy(g1).double() * v.double()).sum())
    rhs1 = float((O.correlation_mean(v.double(), f2.double()) * go.double()).sum())
    lhs2 = float((torch.from_numpy(g2).double() * v.double()).sum())
    rhs2 = float((O.correlation_mean(f1.double(), v.double()) * go.double()).sum())
    assert lhs1 == pytest.approx(rhs1, rel=1e-5) and lhs2 == pytest.approx(rhs2, rel=1e-5)


@pytest.mark.gpu
def test_corr_bwd_kernel_matches_golden(golden_dir):
    from opticalflow_b200.models.correlation_package.correlation import correlation_backward
    g, cases = _cases(golden_dir)
    for i, (lvl, B, H, W, C) in enumerate(cases):
        f1, f2, go = _inputs(lvl, B, H, W, C)
        g1, g2 = correlation_backward(f1.cuda(), f2.cuda(), go.cuda())
        for mine, ref in ((g1, g[f"g1_{i}"]), (g2, g[f"g2_{i}"])):
            assert np.abs(mine.cpu().numpy() - ref).max() <= TOL * np.abs(ref).max()


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 32, 28, 64), (2, 7, 33, 37), (1, 196, 7, 16), (1, 1, 1, 1), (1, 4, 3, 70)])
def test_corr_bwd_kernel_matches_oracle(shape):
    from opticalflow_b200.models.correlation_package.correlation import correlation_backward
    B, C, H, W = shape
    f1, f2, go = _inputs(2, B, H, W, C)
    r1, r2 = O.correlation_backward_numpy(f1.numpy(), f2.numpy(), go.numpy())
    g1, g2 = correlation_backward(f1.cuda(), f2.cuda(), go.cuda())
    assert np.abs(g1.cpu().numpy() - r1).max() <= TOL * max(np.abs(r1).max(), 1e-30)
    assert np.abs(g2.cpu().numpy() - r2).max() <= TOL * max(np.abs(r2).max(), 1e-30)
    only1, none2 = correlation_backward(f1.cuda(), f2.cuda(), go.cuda(), need2=False)
    assert none2 is None and torch.equal(only1, g1)             # deterministic, and each half is optional


@pytest.mark.gpu
def test_correlation_autograd_through_the_module():
    """Correlation(...)(a, b).backward() as the reference's training scripts use it (correlation.py:92-117)."""
    from opticalflow_b200.models.correlation_package.correlation import Correlation
    f1, f2, go = _inputs(5, 2, 10, 18, 12)
    a, b = f1.cuda().requires_grad_(True), f2.cuda().requires_grad_(True)
    out = Correlation(pad_size=4, kernel_size=1, max_displacement=4, stride1=1, stride2=1, corr_multiply=1)(a, b)
    out.backward(go.cuda())
    r1, r2 = O.correlation_backward_numpy(f1.numpy(), f2.numpy(), go.numpy())
    assert np.abs(a.grad.cpu().numpy() - r1).max() <= TOL * np.abs(r1).max()
    assert np.abs(b.grad.cpu().numpy() - r2).max() <= TOL * np.abs(r2).max()
    a2 = f1.cuda().requires_grad_(True)
    Correlation(4, 1, 4, 1, 1)(a2, f2.cuda()).backward(go.cuda())          # only input1 needs a gradient
    assert torch.equal(a2.grad, a.grad)


@pytest.mark.gpu
def test_standalone_warp_matches_reference_golden(golden_dir):
    from opticalflow_b200 import ops
    g = np.load(os.path.join(golden_dir, "warp_corr_op.npz"))
    lvl, B, H, W, C = [int(v) for v in g["meta"]]
    _, f2, flow = O.corr_microbench_inputs(lvl, B, H, W, C)
    # the golden is ATen's CPU grid_sample (products rounded one by one); the kernel accumulates with FMAs in the same
    # corner order as ATen's CUDA grid_sample: a few ulps of the largest value apart
    tol = 1e-6 * float(np.abs(g["warp"]).max())
    w = ops.warp(f2.cuda(), (flow * 1.25).cuda())
    np.testing.assert_allclose(w.cpu().numpy(), g["warp"], rtol=0, atol=tol)
    w2 = ops.warp(f2.cuda(), flow.cuda(), flow_scale=1.25)
    np.testing.assert_allclose(w2.cpu().numpy(), g["warp"], rtol=0, atol=tol)


@pytest.mark.gpu
@pytest.mark.parametrize("align", [False, True])
@pytest.mark.parametrize("shape", [(2, 5, 17, 23), (1, 64, 56, 128), (1, 3, 1, 9)])
def test_standalone_warp_matches_oracle(shape, align):
    from opticalflow_b200 import ops
    from opticalflow_b200.models.PWCNet import PWCDCNet
    B, C, H, W = shape
    gen = torch.Generator().manual_seed(11)
    x = torch.randn(B, C, H, W, generator=gen)
    flo = 3.0 * torch.randn(B, 2, H, W, generator=gen)
    ref = O.warp(x, flo, align_corners=align)
    got = ops.warp(x.cuda(), flo.cuda(), align_corners=align).cpu()
    # a sample whose mask sits within rounding of 0.9999 may flip; everything else agrees to fp32 rounding
    bad = (got - ref).abs() > 1e-5 * ref.abs().max()
    assert bad.float().mean().item() <= 1e-3
    m = PWCDCNet(align_corners=align)
    assert torch.equal(m.warp(x.cuda(), flo.cuda()).cpu(), got)


@pytest.mark.gpu
def test_standalone_warp_rejects_what_the_reference_rejects():
    from opticalflow_b200 import ops
    x = torch.randn(1, 3, 8, 8)
    with pytest.raises(RuntimeError):
        ops.warp(x, torch.zeros(1, 2, 8, 8))                                   # CPU tensors
    with pytest.raises(RuntimeError, match="Float"):
        ops.warp(x.cuda().double(), torch.zeros(1, 2, 8, 8).cuda().double())
