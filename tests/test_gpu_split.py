"""Split-precision convolutions (conv_mode 'mixed'): two 16-bit planes per tensor, three tensor-core products per K step.
Each kernel is compared with torch's fp32 convolution of the SAME fp32 operands; bar: 2e-5 of the output scale (the dropped
x_lo*W_lo term and the second rounding of the planes are ~2^-21; fp32 summation order ~1e-6)."""
import warnings

import pytest
import torch
import torch.nn.functional as F

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu
TOL = 5e-5      # half planes: W_lo of small weights (|W| ~ 1e-2) falls into the subnormal range: ~17-18 bits survive


def _case(cin, cout, N, H, W, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, cin, H, W, generator=g)
    w = torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)
    b = torch.randn(cout, generator=g) * 0.1
    return x, w, b


@pytest.mark.parametrize("fmt", ["f16", "bf16"])
@pytest.mark.parametrize("case", [
    # cin, cout, stride, N, H, W        WM kernel (stride 1) and per-tap kernel (stride 2): pyramid levels 3-6, L4-6 decoders
    (64, 64, 1, 2, 56, 128), (96, 96, 1, 2, 28, 64), (128, 128, 1, 2, 14, 32), (208, 128, 1, 4, 7, 16), (208, 96, 1, 4, 7, 16),
    (181, 128, 1, 2, 28, 64), (597, 32, 1, 1, 28, 64), (629, 64, 1, 2, 14, 32), (81, 128, 1, 8, 7, 16), (40, 32, 1, 1, 9, 21),
    (32, 64, 2, 2, 112, 256), (64, 96, 2, 2, 56, 128), (96, 128, 2, 2, 28, 64), (128, 208, 2, 2, 14, 32), (16, 32, 2, 1, 64, 96),
])
def test_conv3x3_tc_split_matches_fp32(case, fmt):
    from opticalflow_b200 import ops
    cin, cout, stride, N, H, W = case
    x, w, b = _case(cin, cout, N, H, W, 700 + cin + cout)
    ref = F.leaky_relu(F.conv2d(x, w, b, stride=stride, padding=1), 0.1)
    d = torch.device("cuda:0")
    with ops.half_format(fmt):
        out = ops.conv3x3_tc_split(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), stride=stride)
    out = out.permute(0, 3, 1, 2).cpu()
    tol = TOL if fmt == "f16" else 3e-4          # bf16 planes carry 16 mantissa bits together
    assert (out - ref).abs().max().item() <= tol * ref.abs().max().item()


@pytest.mark.parametrize("case", [(64, 64, 1, 2, 24, 40), (245, 96, 1, 1, 14, 32)])
def test_conv3x3_tc_split_half_configurations(case):
    """split input with single output, single input with split output (the boundaries of the mixed plan)"""
    from opticalflow_b200 import ops
    cin, cout, stride, N, H, W = case
    x, w, b = _case(cin, cout, N, H, W, 800 + cin)
    d = torch.device("cuda:0")
    ref = F.leaky_relu(F.conv2d(x, w, b, stride=stride, padding=1), 0.1)
    with ops.half_format("f16"):
        a = ops.conv3x3_tc_split(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), split_in=True, split_out=False)
        xq = x.half().float()
        refq = F.leaky_relu(F.conv2d(xq, w.half().float(), b, stride=stride, padding=1), 0.1)
        c = ops.conv3x3_tc_split(xq.permute(0, 2, 3, 1).contiguous().to(d), w.half().float().to(d), b.to(d), split_in=False, split_out=True)
    s = ref.abs().max().item()
    assert (a.permute(0, 3, 1, 2).cpu() - ref).abs().max().item() <= 1e-3 * s          # output rounded to half
    assert (c.permute(0, 3, 1, 2).cpu() - refq).abs().max().item() <= TOL * s           # half operands, fp32-grade output


@pytest.mark.parametrize("fmt", ["f16", "bf16"])
@pytest.mark.parametrize("case", [(16, 16, 2, 40, 72), (32, 32, 2, 37, 70), (16, 16, 1, 224, 512), (32, 32, 1, 112, 256), (16, 32, 1, 20, 33),
                                  (32, 16, 1, 9, 130)])
@pytest.mark.parametrize("stride", [1, 2])
def test_conv3x3_thin_split_matches_fp32(case, fmt, stride):
    from opticalflow_b200 import ops
    cin, cout, N, H, W = case
    if stride == 2 and cin != 16:
        pytest.skip("stride 2 exists for 16 input channels (conv2a): two 17 x 65 patches of wider pixels exceed shared memory")
    x, w, b = _case(cin, cout, N, H, W, 900 + cin + cout)
    ref = F.leaky_relu(F.conv2d(x, w, b, stride=stride, padding=1), 0.1)
    d = torch.device("cuda:0")
    with ops.half_format(fmt):
        out = ops.conv3x3_thin_split(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), stride=stride)
    out = out.permute(0, 3, 1, 2).cpu()
    tol = TOL if fmt == "f16" else 3e-4
    assert (out - ref).abs().max().item() <= tol * ref.abs().max().item()


@pytest.mark.parametrize("case", [(16, 16, 2, 224, 512), (32, 32, 2, 112, 256), (16, 16, 1, 37, 70), (32, 32, 3, 9, 21), (64, 64, 1, 56, 128),
                                  (533, 32, 1, 28, 64), (565, 64, 2, 14, 32)])
def test_conv3x3_tc_split_tap_lanes(case):
    """split input and output on the tap-lane form of the tcgen05 kernel (pyramid levels 1-2 of the mixed plan; decoder shapes)"""
    from opticalflow_b200 import ops
    cin, cout, N, H, W = case
    x, w, b = _case(cin, cout, N, H, W, 1000 + cin + cout)
    ref = F.leaky_relu(F.conv2d(x, w, b, padding=1), 0.1)
    d = torch.device("cuda:0")
    with ops.half_format("f16"):
        out = ops.conv3x3_tc_split(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), tl=True)
    out = out.permute(0, 3, 1, 2).cpu()
    assert (out - ref).abs().max().item() <= TOL * ref.abs().max().item()
