"""GPU parity tests (run with -m gpu on the B200 box): every CUDA kernel, called through the C ABI, against the
CPU oracle (oracle/pwcnet_oracle.py) on identical seeded inputs.

Tolerances (BASELINE.json north_star): cost volumes 1e-5 relative (here: max|a-b| <= 1e-5 * max|ref|, fp32);
final flow 1e-3 px max-abs in fp32 mode.  Mask flips at the 0.9999 threshold (SURVEY.md hard part 6) are counted
and must be rare.
"""
import warnings

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import pwcnet_oracle as O

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu

LEVELS = {6: (196, 7, 16), 5: (128, 14, 32), 4: (96, 28, 64), 3: (64, 56, 128), 2: (32, 112, 256)}


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _rel_err(a, ref):
    return (a - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)


def _robust_corr_check(out, ref, tol=1e-5, max_bad_frac=2e-4, flow=None, scale=1.0, align_corners=False, mask_eps=2e-6):
    """max-abs relative to the tensor scale.  The only elements allowed beyond `tol` are those that read a warped pixel
    whose mask sum (PWCNet.py:168-174) lies within `mask_eps` of the 0.9999 threshold in the oracle -- there the mask may
    flip between grid_sample's and the fused kernel's fp32 evaluation order; such elements stay below `max_bad_frac`."""
    scale_t = ref.abs().max().item()
    err = (out - ref).abs()
    bad = (err > tol * scale_t)
    frac = bad.float().mean().item()
    assert frac <= max_bad_frac, f"{bad.sum().item()} elements ({frac:.2e}) beyond {tol} relative; max {err.max().item()/scale_t:.3e}"
    if bad.any():
        assert flow is not None, f"{bad.sum().item()} elements beyond {tol} relative without a warp to explain them"
        B, _, H, W = ref.shape
        near = ((O.warp_mask_sum(flow * scale, H, W, align_corners) - 0.9999).abs() <= mask_eps)       # [B,H,W]
        pad = torch.zeros(B, H + 8, W + 8, dtype=torch.bool)
        pad[:, 4:H + 4, 4:W + 4] = near
        for n, d, y, x in bad.nonzero().tolist():
            dy, dx = d // 9 - 4, d % 9 - 4
            assert pad[n, y + dy + 4, x + dx + 4], \
                f"element {(n, d, y, x)} is off by {err[n, d, y, x].item() / scale_t:.2e} and its warped pixel is not at the mask threshold"
    return frac


@pytest.mark.parametrize("shape", [(2, 32, 12, 20), (1, 196, 7, 16), (2, 3, 9, 33), (1, 64, 40, 70)])
def test_correlation_op_nchw(shape):
    """Correlation(...)(f1,f2) == correlation_package semantics (mean over C), incl. ragged shapes."""
    from opticalflow_b200.models.correlation_package.correlation import Correlation
    B, C, H, W = shape
    g = torch.Generator().manual_seed(7)
    f1 = torch.randn(B, C, H, W, generator=g); f2 = torch.randn(B, C, H, W, generator=g)
    ref = O.correlation_mean(f1, f2)
    corr = Correlation(pad_size=4, kernel_size=1, max_displacement=4, stride1=1, stride2=1, corr_multiply=1)
    out = corr(f1.to(_dev()), f2.to(_dev())).cpu()
    assert out.shape == ref.shape
    assert _rel_err(out, ref) <= 1e-5


@pytest.mark.parametrize("lvl", [6, 5, 4, 3, 2])
@pytest.mark.parametrize("layout", ["nchw", "nhwc"])
def test_fused_warp_corr_levels(lvl, layout):
    """All five pyramid-level shapes of the 1024x448 config (batch 2 here), SURVEY 8(d) microbench tensors."""
    from opticalflow_b200 import ops
    C, H, W = LEVELS[lvl]
    f1, f2, flow = O.corr_microbench_inputs(lvl, 2, H, W, C, with_flow=(lvl < 6))
    scale = O.FLOW_SCALE.get(lvl, 1.0)
    ref = O.warp_corr_lrelu(f1, f2, flow, scale)
    d = _dev()
    if layout == "nchw":
        out = ops.warp_corr_lrelu(f1.to(d), f2.to(d), None if flow is None else flow.to(d), scale, layout="nchw").cpu()
    else:
        nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(d)  # noqa: E731
        out = ops.warp_corr_lrelu(nhwc(f1), nhwc(f2), None if flow is None else nhwc(flow), scale, layout="nhwc")
        out = out.permute(0, 3, 1, 2).contiguous().cpu()
    _robust_corr_check(out, ref, flow=flow, scale=scale)


@pytest.mark.parametrize("case", ["zero", "huge", "subpixel", "align_corners"])
def test_fused_warp_corr_edge_flows(case):
    from opticalflow_b200 import ops
    B, C, H, W = 2, 32, 21, 45           # ragged: not multiples of the 8x32 tile
    g = torch.Generator().manual_seed(3)
    f1 = torch.randn(B, C, H, W, generator=g); f2 = torch.randn(B, C, H, W, generator=g)
    ac = False
    if case == "zero":
        flow = torch.zeros(B, 2, H, W)
    elif case == "huge":
        flow = 60.0 * torch.randn(B, 2, H, W, generator=g)    # most samples leave the image, on all four sides
    elif case == "subpixel":
        flow = torch.rand(B, 2, H, W, generator=g) - 0.5
    else:
        flow = 1.5 * torch.randn(B, 2, H, W, generator=g); ac = True
    ref = O.warp_corr_lrelu(f1, f2, flow, 1.25, align_corners=ac)
    d = _dev()
    out = ops.warp_corr_lrelu(f1.to(d), f2.to(d), flow.to(d), 1.25, align_corners=ac, layout="nchw").cpu()
    _robust_corr_check(out, ref, flow=flow, scale=1.25, align_corners=ac)
    if case == "zero":
        # 1-px border of the warped features is masked -> displacement (0,0) is 0 there
        assert out[:, 40, 0, :].abs().max() == 0 and out[:, 40, :, 0].abs().max() == 0


@pytest.mark.parametrize("case", [(2, 32, 30, 72), (1, 64, 56, 128), (2, 96, 28, 64), (2, 196, 7, 16)])
def test_fused_warp_corr_half_features(case):
    """IEEE-half features / output (conv_mode 'f16'): tcgen05 kernel (C <= 64) and warp-MMA kernel (C > 64)"""
    from opticalflow_b200 import ops
    B, C, H, W = case
    f1, f2, flow = O.corr_microbench_inputs(3, B, H, W, C)
    f1h, f2h = f1.half(), f2.half()
    ref = O.warp_corr_lrelu(f1h.float(), f2h.float(), flow, 2.5)
    d = _dev()
    nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(d)  # noqa: E731
    Cp = (C + 15) // 16 * 16
    a, b = torch.zeros(B, H, W, Cp, dtype=torch.float16, device=d), torch.zeros(B, H, W, Cp, dtype=torch.float16, device=d)
    a[..., :C], b[..., :C] = nhwc(f1h), nhwc(f2h)
    out = torch.zeros(B, H, W, 88, dtype=torch.float16, device=d)
    with ops.half_format("f16"):
        ops.warp_corr_lrelu(a[..., :C], b[..., :C], nhwc(flow), 2.5, layout="nhwc", out=out[..., :81])
    out = out[..., :81].float().permute(0, 3, 1, 2).cpu()
    assert (out - ref).abs().max().item() <= 1.5e-3 * ref.abs().max().item()    # half rounding of output and warped operand


def test_fused_warp_corr_bf16_features():
    from opticalflow_b200 import ops
    C, H, W = 64, 24, 40
    f1, f2, flow = O.corr_microbench_inputs(3, 2, H, W, C)
    f1b, f2b = f1.bfloat16(), f2.bfloat16()
    ref = O.warp_corr_lrelu(f1b.float(), f2b.float(), flow, 2.5)
    d = _dev()
    nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(d)  # noqa: E731
    out = ops.warp_corr_lrelu(nhwc(f1b), nhwc(f2b), nhwc(flow), 2.5, layout="nhwc")
    assert out.dtype == torch.bfloat16
    out = out.float().permute(0, 3, 1, 2).cpu()
    assert (out - ref).abs().max().item() <= 8e-3 * ref.abs().max().item()    # bf16 output rounding (2^-8)


@pytest.mark.parametrize("case", [
    # B, C, H, W, with_flow        (C <= 64 bf16 NHWC -> tensor-core cost volume, csrc/warp_corr_tc.cuh)
    (2, 32, 112, 256, True), (1, 64, 56, 128, True), (2, 32, 30, 72, True), (1, 16, 5, 24, True), (1, 48, 9, 33, True),
    (2, 64, 19, 40, False), (3, 32, 64, 96, True),
    # C > 64: small-level kernel (csrc/warp_corr_small.cuh), the level 4/5/6 shapes and a ragged one
    (2, 96, 28, 64, True), (2, 128, 14, 32, True), (2, 196, 7, 16, False), (1, 100, 9, 45, True),
    # image-2 features with the pyramid buffers' padded pitch (C rounded up to 16): warp-level MMA kernel
    # (csrc/warp_corr_mma.cuh) at every width, incl. the half channel group of C = 196 / 100 and 16-pixel strips
    (2, 196, 7, 16, False, 16), (2, 196, 7, 16, True, 16), (1, 100, 9, 45, True, 16), (2, 128, 14, 32, True, 16),
    (1, 96, 28, 64, True, 16), (1, 32, 37, 50, True, 16), (1, 64, 11, 16, True, 16), (8, 32, 112, 256, True, 16),
])
def test_fused_warp_corr_tensor_core_path(case):
    """bf16 features inside concat-buffer slices (pitch 544+C: f1 at channel 544, cost volume at 448), ragged shapes, more
    items than one wave.  Bar: bf16 rounding of the output plus of the warped operand (both 2^-9 relative)."""
    from opticalflow_b200 import ops
    B, C, H, W, with_flow = case[:5]
    f2_round = case[5] if len(case) > 5 else 1
    f1, f2, flow = O.corr_microbench_inputs(2, B, H, W, C, with_flow=with_flow)
    f1b, f2b = f1.bfloat16(), f2.bfloat16()
    ref = O.warp_corr_lrelu(f1b.float(), f2b.float(), flow, 5.0)
    d = _dev()
    cat = torch.full((B, H, W, 544 + 32 * ((C + 31) // 32)), 7.0, device=d, dtype=torch.bfloat16)
    cat[..., 544:544 + C] = f1b.permute(0, 2, 3, 1).to(d)
    f2n = torch.full((B, H, W, (C + f2_round - 1) // f2_round * f2_round), 3.0, device=d, dtype=torch.bfloat16)[..., :C]
    f2n.copy_(f2b.permute(0, 2, 3, 1))
    fl = None if flow is None else flow.permute(0, 2, 3, 1).contiguous().to(d)
    ops.warp_corr_lrelu(cat[..., 544:544 + C], f2n, fl, 5.0, layout="nhwc", out=cat[..., 448:529])
    torch.cuda.synchronize()
    out = cat[..., 448:529].float().permute(0, 3, 1, 2).cpu()
    scale = ref.abs().max().item()
    err = (out - ref).abs()
    assert err.max().item() <= 8e-3 * scale, f"max err {err.max().item() / scale:.3e} of scale"
    assert err.mean().item() <= 1e-3 * scale
    # neighbours of the slice are untouched
    assert (cat[..., :448] == 7.0).all() and (cat[..., 529:544] == 7.0).all()


CONV_CASES = [
    # cin, cout, stride, dil, H, W
    (4, 16, 2, 1, 32, 48), (16, 16, 1, 1, 17, 23), (32, 64, 2, 1, 16, 32), (96, 128, 2, 1, 14, 18),
    (196, 196, 1, 1, 7, 16), (128, 96, 1, 8, 20, 36), (96, 64, 1, 16, 24, 40), (128, 128, 1, 2, 9, 31),
    (32, 2, 1, 1, 12, 20), (64, 32, 1, 1, 8, 16),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv3x3_f32_matches_torch(case):
    from opticalflow_b200 import ops
    cin, cout, stride, dil, H, W = case
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, cin, H, W, generator=g)
    w = torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)
    b = torch.randn(cout, generator=g) * 0.1
    slope = 1.0 if cout == 2 else 0.1
    ref = F.leaky_relu(F.conv2d(x, w, b, stride=stride, padding=dil, dilation=dil), slope)
    d = _dev()
    out = ops.conv3x3_f32(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), stride, dil, slope)
    out = out.permute(0, 3, 1, 2).cpu()
    assert out.shape == ref.shape
    assert (out - ref).abs().max().item() <= 2e-5 * max(ref.abs().max().item(), 1.0)


def test_conv3x3_f32_concat_slice_and_chan_map():
    """Reads a channel suffix of a wide buffer through chan_map and writes the slice in front."""
    from opticalflow_b200 import ops, engine
    lvl, H, W = 4, 12, 20
    wd = engine.concat_width(lvl)
    k = 2                                   # conv4_2: reads [192, end), writes [96,192)
    start, wr = engine.DEC_START[k], engine.DEC_WRITE[k]
    cm = engine.chan_map(lvl, start)
    cin = len(cm)
    g = torch.Generator().manual_seed(5)
    buf = torch.randn(1, H, W, wd, generator=g)
    buf[..., 533:544] = 0
    w = torch.randn(96, cin, 3, 3, generator=g) / (3 * cin ** 0.5)
    b = torch.randn(96, generator=g) * 0.1
    xin = buf[..., start:][..., cm].permute(0, 3, 1, 2).contiguous()     # torch channel order
    ref = F.leaky_relu(F.conv2d(xin, w, b, padding=1), 0.1)
    d = _dev()
    dbuf = buf.to(d)
    before = dbuf.clone()
    ops.conv3x3_f32(dbuf, w.to(d), b.to(d), 1, 1, 0.1, in_coff=start, cin_span=wd - start, out=dbuf, out_coff=wr, chan_map=cm)
    got = dbuf[..., wr:wr + 96].permute(0, 3, 1, 2).cpu()
    assert (got - ref).abs().max().item() <= 2e-5 * max(ref.abs().max().item(), 1.0)
    keep = torch.ones(wd, dtype=torch.bool); keep[wr:wr + 96] = False
    assert torch.equal(dbuf[..., keep], before[..., keep])              # nothing else was touched


@pytest.mark.parametrize("cin", [2, 597])
def test_deconv4x4s2_matches_torch(cin):
    from opticalflow_b200 import ops
    g = torch.Generator().manual_seed(13)
    H, W = 9, 14
    pitch = (cin + 3) // 4 * 4
    x = torch.zeros(2, H, W, pitch)
    x[..., :cin] = torch.randn(2, H, W, cin, generator=g)
    w = torch.randn(cin, 2, 4, 4, generator=g) / (4 * cin ** 0.5)
    b = torch.randn(2, generator=g) * 0.1
    ref = F.conv_transpose2d(x[..., :cin].permute(0, 3, 1, 2), w, b, stride=2, padding=1)
    d = _dev()
    out = ops.deconv4x4s2_c2_f32(x.to(d), w.to(d), b.to(d)).permute(0, 3, 1, 2).cpu()
    assert (out - ref).abs().max().item() <= 2e-5 * max(ref.abs().max().item(), 1.0)
