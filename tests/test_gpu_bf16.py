"""GPU parity of the bf16-conv mode: tcgen05 convolutions, bf16 companions, and the whole forward against the CPU
oracle.  Bars (BASELINE.json north_star): mean EPE delta <= 0.02 px for the bf16 conv mode; single kernels are
compared with torch fp32 run on the bf16-rounded operands (tolerance = bf16 output rounding, 2^-8 relative)."""
import warnings

import pytest
import torch
import torch.nn.functional as F

from oracle import pwcnet_oracle as O

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu

TC_CASES = [
    # cin, cout, stride, dil, N, H, W, out_f32
    (64, 128, 1, 1, 2, 24, 48, 0), (128, 128, 1, 1, 2, 21, 37, 0), (32, 32, 1, 1, 2, 16, 32, 0),
    (16, 16, 1, 1, 2, 16, 32, 0), (96, 64, 1, 16, 1, 40, 64, 0), (128, 96, 1, 8, 1, 24, 40, 0),
    (576, 128, 1, 1, 1, 16, 32, 0), (208, 208, 1, 1, 2, 7, 16, 0), (576, 2, 1, 1, 1, 16, 32, 1),
    (32, 2, 1, 1, 2, 16, 32, 1), (480, 64, 1, 1, 1, 16, 16, 0), (64, 96, 2, 1, 2, 32, 64, 0), (16, 32, 2, 1, 1, 32, 32, 0),
    # dilation 16: y-dense x-split mode (W % 16 == 0), its fallback (ragged W), dilation 8 y-dense with ragged H and W
    (64, 64, 1, 16, 2, 50, 96, 0), (96, 64, 1, 16, 1, 40, 70, 0), (128, 96, 1, 8, 2, 61, 43, 0), (96, 32, 1, 16, 1, 112, 256, 0),
]


@pytest.mark.parametrize("case", [
    # cin, cout, N, H, W, residual      tap-stacked mode (9*cout <= 128): predict_flowL / pfupL / dc_conv7 shapes
    (576, 2, 1, 16, 32, False), (32, 2, 2, 16, 32, True), (565, 2, 2, 112, 256, False), (629, 10, 1, 56, 128, False),
    (597, 10, 2, 7, 16, False), (80, 14, 1, 23, 41, False), (64, 2, 3, 50, 9, True),
])
def test_conv3x3_tc_tap_stacked(case):
    from opticalflow_b200 import ops
    cin, cout, N, H, W, with_res = case
    g = torch.Generator().manual_seed(300 + cin + cout)
    x = torch.randn(N, cin, H, W, generator=g).bfloat16()
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).bfloat16().float()
    b = torch.randn(cout, generator=g) * 0.1
    res = torch.randn(N, H, W, 4, generator=g) if with_res else None
    ref = F.conv2d(x.float(), w, b, padding=1)
    if with_res:
        ref = ref + res.permute(0, 3, 1, 2)[:, :cout]
    d = torch.device("cuda:0")
    span = (cin + 15) // 16 * 16
    xn = torch.zeros(N, H, W, span, dtype=torch.bfloat16)
    xn[..., :cin] = x.permute(0, 2, 3, 1)
    out = ops.conv3x3_tc(xn.to(d), w.to(d), b.to(d), 1, 1, 1.0, cin_span=span, out_f32=True, ts=True,
                         residual=None if res is None else res.to(d))
    out = out.float().permute(0, 3, 1, 2).cpu()[:, :cout]
    assert (out - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


@pytest.mark.parametrize("fmt", ["bf16", "f16"])
@pytest.mark.parametrize("case", [
    # cin, cout, N, H, W      tap-lane mode (Cout <= 64): decoder conv{2,3}_3 / _4 and dc_conv6 shapes, ragged tiles, odd channel counts
    (533, 32, 2, 112, 256), (469, 64, 1, 112, 256), (565, 32, 1, 56, 128), (501, 64, 2, 56, 128), (64, 32, 1, 112, 256),
    (128, 64, 2, 21, 37), (96, 32, 1, 7, 16), (72, 48, 1, 30, 9), (192, 20, 3, 13, 50), (64, 64, 8, 14, 32),
])
def test_conv3x3_tc_tap_lanes(case, fmt):
    from opticalflow_b200 import ops
    cin, cout, N, H, W = case
    dt = torch.float16 if fmt == "f16" else torch.bfloat16
    g = torch.Generator().manual_seed(400 + cin + cout)
    x = torch.randn(N, cin, H, W, generator=g).to(dt)
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).to(dt).float()
    b = torch.randn(cout, generator=g) * 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w, b, padding=1), 0.1)
    d = torch.device("cuda:0")
    span = (cin + 15) // 16 * 16
    xn = torch.zeros(N, H, W, span, dtype=dt)
    xn[..., :cin] = x.permute(0, 2, 3, 1)
    with ops.half_format(fmt):
        out = ops.conv3x3_tc(xn.to(d), w.to(d), b.to(d), 1, 1, 0.1, cin_span=span, tl=True)
    got = out.float().permute(0, 3, 1, 2).cpu()
    assert (got[:, :cout] - ref).abs().max().item() <= (1.5e-3 if fmt == "f16" else 1e-2) * ref.abs().max().item()
    assert got[:, cout:].abs().max().item() == 0.0 if got.shape[1] > cout else True


@pytest.mark.parametrize("fmt", ["bf16", "f16"])
@pytest.mark.parametrize("case", [
    # cin, cout, nhead, N, H, W      conv2_4 + predict_flow2 / conv3_4 + predict_flow3|upfeat3 shapes, ragged tiles
    (533, 32, 2, 2, 112, 256), (565, 32, 10, 1, 56, 128), (96, 32, 2, 1, 7, 16), (192, 20, 10, 3, 13, 50), (64, 16, 5, 2, 30, 9),
])
def test_conv3x3_tc_head_rows_and_partial(case, fmt):
    """Two layers over one input in one pass (pwc_tc_conv_set_head_rows): the tap-lane layer's own output is unchanged, the raw sums
    of the extra channels equal a bias-free conv, and a tap-stacked layer over 32 further channels + that partial
    (pwc_tc_conv_set_partial) equals the head over all channels."""
    from opticalflow_b200 import ops
    cin, cout, nh, N, H, W = case
    dt = torch.float16 if fmt == "f16" else torch.bfloat16
    g = torch.Generator().manual_seed(700 + cin + cout + nh)
    x = torch.randn(N, cin, H, W, generator=g).to(dt)
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).to(dt).float()
    b = torch.randn(cout, generator=g) * 0.1
    wh = (torch.randn(nh, cin + 32, 3, 3, generator=g) / (3 * (cin + 32) ** 0.5)).to(dt).float()     # head over [32 new | cin]
    bh = torch.randn(nh, generator=g) * 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w, b, padding=1), 0.1)
    ref_part = F.conv2d(x.float(), wh[:, 32:], None, padding=1)
    d = torch.device("cuda:0")
    span = (cin + 15) // 16 * 16
    xn = torch.zeros(N, H, W, span, dtype=dt)
    xn[..., :cin] = x.permute(0, 2, 3, 1)
    with ops.half_format(fmt):
        out = ops.conv3x3_tc(xn.to(d), w.to(d), b.to(d), 1, 1, 0.1, cin_span=span, tl=True, head_w=wh[:, 32:].to(d))
        part = ops.conv3x3_tc.head
        got = out.float().permute(0, 3, 1, 2).cpu()
        tol = 1.5e-3 if fmt == "f16" else 1e-2
        assert (got[:, :cout] - ref).abs().max().item() <= tol * ref.abs().max().item()
        gp = part.permute(0, 3, 1, 2).cpu()
        assert (gp - ref_part).abs().max().item() <= 2e-5 * max(1.0, ref_part.abs().max().item())       # fp32 sums of exact products
        # the head: 32 new channels (rounded to the 16-bit format as a real layer output would be) + the partial
        y = torch.randn(N, 32, H, W, generator=g).to(dt)
        ref_head = F.conv2d(torch.cat([y.float(), x.float()], 1), wh, bh, padding=1)
        yn = y.permute(0, 2, 3, 1).contiguous().to(d)
        head = ops.conv3x3_tc(yn, wh[:, :32].to(d), bh.to(d), 1, 1, 1.0, cin_span=32, out_f32=True, ts=True, partial=part)
    gh = head.permute(0, 3, 1, 2).cpu()[:, :nh]
    assert (gh - ref_head).abs().max().item() <= 5e-5 * max(1.0, ref_head.abs().max().item())


@pytest.mark.parametrize("case", TC_CASES)
def test_conv3x3_tc_matches_torch(case):
    from opticalflow_b200 import ops
    cin, cout, stride, dil, N, H, W, of32 = case
    g = torch.Generator().manual_seed(100 + cin + cout)
    x = torch.randn(N, cin, H, W, generator=g).bfloat16()
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).bfloat16().float()
    b = torch.randn(cout, generator=g) * 0.1
    slope = 1.0 if of32 else 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w, b, stride=stride, padding=dil, dilation=dil), slope)
    d = torch.device("cuda:0")
    out = ops.conv3x3_tc(x.permute(0, 2, 3, 1).contiguous().to(d), w.to(d), b.to(d), stride, dil, slope, out_f32=bool(of32))
    out = out.float().permute(0, 3, 1, 2).cpu()[:, :cout]
    tol = (2e-3 if of32 else 1e-2) * ref.abs().max().item()
    assert (out - ref).abs().max().item() <= tol


@pytest.mark.parametrize("case", [(64, 128, 1, 1, 2, 24, 48), (576, 128, 1, 1, 1, 16, 32), (96, 64, 1, 16, 1, 40, 64), (64, 96, 2, 1, 2, 32, 64),
                                  (245, 128, 1, 1, 8, 7, 16)])
def test_conv3x3_tc_half_format(case):
    """the same tcgen05 kernels with IEEE-half operands / outputs (pwc_set_half_format(1), conv_mode 'f16'): tolerance = half
    output rounding (2^-11 relative)"""
    from opticalflow_b200 import ops
    cin, cout, stride, dil, N, H, W = case
    g = torch.Generator().manual_seed(500 + cin + cout)
    x = torch.randn(N, cin, H, W, generator=g).half()
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).half().float()
    b = torch.randn(cout, generator=g) * 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w, b, stride=stride, padding=dil, dilation=dil), 0.1)
    d = torch.device("cuda:0")
    span = (cin + 15) // 16 * 16
    xn = torch.zeros(N, H, W, span, dtype=torch.float16)
    xn[..., :cin] = x.permute(0, 2, 3, 1)
    with ops.half_format("f16"):
        out = ops.conv3x3_tc(xn.to(d), w.to(d), b.to(d), stride, dil, 0.1, cin_span=span)
    assert out.dtype == torch.float16
    out = out.float().permute(0, 3, 1, 2).cpu()[:, :cout]
    assert (out - ref).abs().max().item() <= 1.5e-3 * ref.abs().max().item()


@pytest.mark.parametrize("shape", [(2, 64, 96), (1, 72, 200), (2, 130, 516), (1, 448, 1024), (1, 10, 66), (1, 2, 4)])
def test_conv1a_bf16_matches_torch(shape):
    """tiles of 8 x 64 output pixels: ragged in both directions, one tile, many tiles; W % 4 != 0 takes the CUDA-core kernel"""
    from opticalflow_b200 import ops
    g = torch.Generator().manual_seed(5)
    x = torch.rand(shape[0], 6, shape[1], shape[2], generator=g)
    w = torch.randn(16, 3, 3, 3, generator=g) / 5.0
    b = torch.randn(16, generator=g) * 0.1
    d = torch.device("cuda:0")
    out = ops.conv1a_bf16(x.to(d), w.to(d), b.to(d)).float().permute(0, 3, 1, 2).cpu()
    ref = torch.cat([F.leaky_relu(F.conv2d(x[:, :3], w, b, stride=2, padding=1), 0.1),
                     F.leaky_relu(F.conv2d(x[:, 3:], w, b, stride=2, padding=1), 0.1)], 0)
    assert out.shape == ref.shape
    assert (out - ref).abs().max().item() <= 8e-3 * ref.abs().max().item()


def test_conv1a_split_keeps_fp32_grade_inputs():
    """The tensor-core conv1a splits x and w into bf16 hi + lo parts (3 products, lo*lo dropped: ~2^-16 relative error
    against 2^-9 for a plain bf16 cast): its bf16 output must be the correctly rounded fp64 result in >= 99% of the elements
    (measured 99.4%: the rest sit within 2^-16 of a rounding boundary) and never more than one bf16 ulp (+ 2^-15 of the
    sum of |x*w|, for results that cancel to nearly zero) away."""
    from opticalflow_b200 import ops
    g = torch.Generator().manual_seed(6)
    x = torch.rand(2, 6, 96, 256, generator=g)
    w = torch.randn(16, 3, 3, 3, generator=g) / 5.0
    b = torch.randn(16, generator=g) * 0.1
    d = torch.device("cuda:0")
    out = ops.conv1a_bf16(x.to(d), w.to(d), b.to(d)).permute(0, 3, 1, 2).cpu()
    ref64 = torch.cat([F.leaky_relu(F.conv2d(x[:, :3].double(), w.double(), b.double(), stride=2, padding=1), 0.1),
                       F.leaky_relu(F.conv2d(x[:, 3:].double(), w.double(), b.double(), stride=2, padding=1), 0.1)], 0)
    want = ref64.float().bfloat16()
    exact = (out == want).float().mean().item()
    assert exact >= 0.99, exact
    # never more than one bf16 ulp of the value plus the split's own error, 2^-16 of the sum of |x*w| (which matters only
    # where the 27 products cancel to a result near zero)
    mag = torch.cat([F.conv2d(x[:, :3], w.abs(), None, stride=2, padding=1), F.conv2d(x[:, 3:], w.abs(), None, stride=2, padding=1)], 0)
    tol = want.float().abs() * 2.0 ** -7 + mag * 2.0 ** -15
    assert ((out.float() - want.float()).abs() <= tol).all()
    cast = torch.cat([F.leaky_relu(F.conv2d(x[:, :3].bfloat16().float(), w.bfloat16().float(), b, stride=2, padding=1), 0.1),
                      F.leaky_relu(F.conv2d(x[:, 3:].bfloat16().float(), w.bfloat16().float(), b, stride=2, padding=1), 0.1)], 0)
    assert (cast.bfloat16() == want).float().mean().item() < exact - 0.1            # what the split buys


@pytest.mark.parametrize("idt,odt", [(torch.bfloat16, torch.bfloat16), (torch.float32, torch.bfloat16), (torch.bfloat16, torch.float32)])
def test_deconv_dtypes(idt, odt):
    from opticalflow_b200 import ops
    g = torch.Generator().manual_seed(13)
    cin, H, W = 100, 9, 14
    x = torch.randn(2, H, W, cin, generator=g).to(idt)
    w = torch.randn(cin, 2, 4, 4, generator=g) / (4 * cin ** 0.5)
    b = torch.randn(2, generator=g) * 0.1
    ref = F.conv_transpose2d(x.float().permute(0, 3, 1, 2), w, b, stride=2, padding=1)
    d = torch.device("cuda:0")
    out, outb = ops.deconv4x4s2_c2(x.to(d), w.to(d), b.to(d), out_dtype=odt)
    sc = max(ref.abs().max().item(), 1.0)
    assert (outb.permute(0, 3, 1, 2).cpu() - ref).abs().max().item() <= 2e-5 * sc       # fp32 copy is exact-ish
    assert (out.float().permute(0, 3, 1, 2).cpu() - ref).abs().max().item() <= (8e-3 if odt == torch.bfloat16 else 2e-5) * sc


def _model(wseed, rbias, **kw):
    from opticalflow_b200.models.PWCNet import PWCDCNet
    torch.manual_seed(wseed)
    m = PWCDCNet(**kw).eval()
    if rbias:
        m.load_state_dict(O.randomize_biases(m.state_dict()))
    return m.cuda()


def _epe(a, b):
    return (a - b).pow(2).sum(1).sqrt()


@pytest.mark.parametrize("shape,shift", [((1, 128, 192), 1), ((2, 64, 128), 0), ((1, 448, 1024), 0), ((1, 448, 1024), 1)])
def test_forward_mixed_epe_delta(shape, shift):
    """conv_mode 'mixed' (split-precision pyramid / level 4-6 decoders / up_flow + up_feat, half elsewhere): the north_star
    bar in PIXELS -- mean EPE delta <= 0.02 px on the x20 flow (script_pwc.py:72) against the fp32 reference forward, both
    synthetic inputs of SURVEY 8(d) at 1024x448."""
    B, H, W = shape
    m = _model(0, False, conv_mode="mixed")
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    x = O.shifted_pair_input(B, H, W, 3) if shift else O.synthetic_input(B, H, W, 1)
    ref = O.forward(sd, x)
    flow = m(x.cuda()).cpu()
    e_px = 20 * _epe(flow, ref).mean().item()
    print(f"mixed mode {W}x{H} shift={shift}: mean EPE delta vs fp32 reference {e_px:.4f} px (x20 flow), max-abs "
          f"{20 * (flow - ref).abs().max().item():.3f} px; |flow| max {20 * ref.abs().max().item():.1f} px")
    assert e_px <= 0.02
    assert torch.equal(m(x.cuda()).cpu(), flow)


@pytest.mark.parametrize("shape,shift", [((1, 128, 192), 1), ((1, 448, 1024), 0), ((1, 448, 1024), 1)])
def test_forward_f16_epe_delta(shape, shift):
    """conv_mode 'f16' (IEEE-half storage, same kernels): the EPE delta in PIXELS of the x20 flow (script_pwc.py:72) must
    stay below 0.09 px (the storage type's measured floor is 0.04-0.05 px at 1024x448, tools/error_budget.py) and within
    1.5x of what an fp32 forward with half-rounded weights / stored activations gives (oracle storage='f16')."""
    B, H, W = shape
    m = _model(0, False, conv_mode="f16")
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    x = O.shifted_pair_input(B, H, W, 3) if shift else O.synthetic_input(B, H, W, 1)
    ref = O.forward(sd, x)
    sim = O.forward(sd, x, storage="f16")
    flow = m(x.cuda()).cpu()
    e_ref, e_simref = 20 * _epe(flow, ref).mean().item(), 20 * _epe(sim, ref).mean().item()
    print(f"f16 mode {W}x{H} shift={shift}: mean EPE delta vs fp32 reference {e_ref:.4f} px (x20 flow); half-storage oracle vs "
          f"fp32 {e_simref:.4f} px; |flow| max {20 * ref.abs().max().item():.1f} px")
    assert e_ref <= 0.09
    assert e_ref <= 1.5 * e_simref + 2e-3
    assert torch.equal(m(x.cuda()).cpu(), flow)


@pytest.mark.parametrize("shape,shift", [((1, 128, 192), 1), ((1, 448, 1024), 0), ((1, 448, 1024), 1)])
def test_forward_bf16_epe_delta(shape, shift):
    """north_star bar: mean EPE delta <= 0.02 against the fp32 reference forward, reference random init (kaiming
    weights, zero biases: PWCNet.py:134-138) on the synthetic inputs of SURVEY 8(d).

    Units: the EPE is taken on the network output flow2 exactly as PWCDCNet.forward returns it -- the quantity the
    reference's own parity script compares (onnx_pth_compare.py:103-111 feeds the raw model output to its EPE,
    onnx_pth_compare.py:162-167).  The same number in pixels of the x20-scaled flow (script_pwc.py:72) is printed.
    A second bar ties the CUDA path to the precision model: it must be as close to the fp32 reference as an fp32
    forward with bf16-rounded weights / stored activations (oracle bf16_storage=True) is, within 1.5x."""
    B, H, W = shape
    m = _model(0, False, conv_mode="bf16")
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    x = O.shifted_pair_input(B, H, W, 3) if shift else O.synthetic_input(B, H, W, 1)
    ref = O.forward(sd, x)
    sim = O.forward(sd, x, bf16_storage=True)
    flow = m(x.cuda()).cpu()
    e_ref, e_sim, e_simref = _epe(flow, ref).mean().item(), _epe(flow, sim).mean().item(), _epe(sim, ref).mean().item()
    print(f"bf16 mode {W}x{H} shift={shift}: mean EPE delta vs fp32 reference {e_ref:.5f} (x20: {20 * e_ref:.4f} px); "
          f"vs bf16-storage oracle {e_sim:.5f}; bf16-storage oracle vs fp32 {e_simref:.5f}; "
          f"|flow2| max {ref.abs().max().item():.3f} mean {ref.abs().mean().item():.3f}")
    assert e_ref <= 0.02
    assert e_ref <= 1.5 * e_simref + 1e-4
    again = m(x.cuda()).cpu()                 # CUDA-graph replay: bit-identical
    assert torch.equal(again, flow)


def test_forward_bf16_random_bias_tracks_bf16_oracle():
    """Non-zero biases (a bias bug would hide behind the reference's zero init): the flow is ~40 px and chaotic, so
    the bar is relative: the CUDA path must be as close to the fp32 reference as the bf16-storage oracle is (x2)."""
    B, H, W = 1, 128, 192
    m = _model(0, True, conv_mode="bf16")
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    x = O.shifted_pair_input(B, H, W, 3)
    ref = 20.0 * O.forward(sd, x)
    sim = 20.0 * O.forward(sd, x, bf16_storage=True)
    flow = 20.0 * m(x.cuda()).cpu()
    e_ref, e_simref = _epe(flow, ref).mean().item(), _epe(sim, ref).mean().item()
    print(f"random-bias: ours vs fp32 {e_ref:.4f} px, bf16-storage oracle vs fp32 {e_simref:.4f} px, absmax {ref.abs().max().item():.1f}")
    assert e_ref <= 2.0 * e_simref + 1e-3


@pytest.mark.parametrize("case", [
    # cin, cout, stride, N, H, W          (levels 5 / 6 of 1024x448 and ragged shapes; coutp = 208 for 196)
    (81, 128, 1, 8, 7, 16), (565, 32, 1, 8, 7, 16), (196, 196, 1, 16, 7, 16), (213, 128, 1, 8, 14, 32),
    (469, 64, 1, 2, 14, 32), (128, 196, 2, 4, 14, 32), (96, 128, 2, 2, 28, 64), (40, 24, 1, 3, 9, 21), (72, 48, 2, 1, 11, 19),
])
def test_conv3x3_mma_small_matches_torch(case):
    """pwc_conv3x3_mma_fwd: every tile configuration the heuristic can pick plus the forced ones, stride 1 and 2."""
    import os
    from opticalflow_b200 import ops
    cin, cout, stride, N, H, W = case
    g = torch.Generator().manual_seed(cin * 7 + cout)
    x = (torch.randn(N, cin, H, W, generator=g)).bfloat16()
    w = torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)
    b = torch.randn(cout, generator=g) * 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w.bfloat16().float(), b, stride=stride, padding=1), 0.1)
    d = torch.device("cuda:0")
    pitch = (cin + 7) // 8 * 8 + 8
    xn = torch.full((N, H, W, pitch), 3.0, dtype=torch.bfloat16)           # neighbours of the slice must not leak in
    xn[..., :cin] = x.permute(0, 2, 3, 1)
    xn = xn.to(d)
    out = ops.conv3x3_mma_small(xn, w.to(d), b.to(d), stride, 0.1, cin_span=cin)
    coutp = out.shape[-1]
    o = out.float().permute(0, 3, 1, 2).cpu()
    tol = 1e-2 * ref.abs().max().item()
    assert (o[:, :cout] - ref).abs().max().item() <= tol
    assert (o[:, cout:] == 0).all()                                        # padded output channels are written as zeros
    # two inputs (im1 / im2 feature tensors of the stride-2 pyramid layers)
    if N % 2 == 0:
        h = N // 2
        out2 = ops.conv3x3_mma_small(xn[:h].contiguous(), w.to(d), b.to(d), stride, 0.1, cin_span=cin,
                                     x2=xn[h:].contiguous(), n_split_in=h)
        assert torch.equal(out2, out)


@pytest.mark.parametrize("case", [
    # cin, cout, stride, N, H, W     (the 8 pyramid layers' channel pairs; ragged tiles; more tiles than resident CTAs)
    (16, 16, 1, 2, 24, 40), (16, 32, 2, 2, 30, 70), (32, 32, 1, 1, 9, 33), (16, 64, 2, 3, 17, 31), (64, 64, 1, 2, 16, 64),
    (16, 16, 1, 4, 224, 512), (16, 32, 2, 2, 224, 512), (64, 32, 1, 1, 8, 32), (16, 16, 2, 1, 5, 7), (32, 64, 1, 2, 56, 128),
    (16, 16, 1, 3, 13, 100), (16, 16, 1, 1, 1, 1), (16, 16, 1, 2, 9, 129),      # the 16 -> 16 kernel: ragged rows / columns
    (32, 32, 1, 2, 24, 70), (32, 32, 1, 4, 112, 256), (32, 32, 1, 1, 1, 1),     # the 32 -> 32 kernel
])
def test_conv3x3_thin_matches_torch(case):
    """pwc_conv3x3_thin_fwd against torch conv2d on the same bf16-rounded operands; channel slices of wider tensors."""
    from opticalflow_b200 import ops
    cin, cout, stride, N, H, W = case
    g = torch.Generator().manual_seed(cin * 3 + cout + stride)
    x = torch.randn(N, cin, H, W, generator=g).bfloat16()
    w = torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)
    b = torch.randn(cout, generator=g) * 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w.bfloat16().float(), b, stride=stride, padding=1), 0.1)
    d = torch.device("cuda:0")
    xn = torch.full((N, H, W, cin + 16), 3.0, dtype=torch.bfloat16)           # input slice at channel offset 8
    xn[..., 8:8 + cin] = x.permute(0, 2, 3, 1)
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    out = torch.full((N, ho, wo, cout + 24), 7.0, device=d, dtype=torch.bfloat16)   # output slice at channel offset 16
    ops.conv3x3_thin(xn.to(d), w.to(d), b.to(d), stride, 0.1, in_coff=8, out=out, out_coff=16)
    o = out[..., 16:16 + cout].float().permute(0, 3, 1, 2).cpu()
    assert (o - ref).abs().max().item() <= 1e-2 * ref.abs().max().item()
    assert (out[..., :16] == 7.0).all() and (out[..., 16 + cout:] == 7.0).all()  # neighbours of the slice untouched
    if N % 2 == 0:                                                               # im1 | im2 halves from two tensors
        h = N // 2
        xd = xn.to(d)
        out2 = torch.full_like(out, 7.0)
        ops.conv3x3_thin(xd[:h].contiguous(), w.to(d), b.to(d), stride, 0.1, in_coff=8, x2=xd[h:].contiguous(), n_split_in=h,
                         out=out2, out_coff=16)
        assert torch.equal(out2, out)


def test_conv3x3_thin_rejects_what_it_cannot_stage():
    """stride 2 with Cin >= 32 needs two 17 x 65 patches of > 88 KB: the entry point returns an error code (no fallback)."""
    from opticalflow_b200 import ops
    d = torch.device("cuda:0")
    x = torch.zeros(1, 16, 32, 32, device=d, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        ops.conv3x3_thin(x, torch.zeros(64, 32, 3, 3, device=d), torch.zeros(64, device=d), 2, 0.1)
    with pytest.raises(RuntimeError):
        ops.conv3x3_thin(torch.zeros(1, 8, 8, 24, device=d, dtype=torch.bfloat16), torch.zeros(16, 24, 3, 3, device=d),
                         torch.zeros(16, device=d), 1, 0.1)
