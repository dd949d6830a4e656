"""Functional wrappers over the C ABI for tests, the microbenchmark and callers that hold torch tensors."""
import torch

from . import _lib


class half_format:
    """``with ops.half_format('f16'):`` -- the 16-bit tensors handed to the wrappers below are IEEE half instead of bf16
    (pwc_set_half_format: thread-local, read by the library at create / launch time)."""

    def __init__(self, fmt):
        self.fmt = {"bf16": 0, "f16": 1}[fmt]

    def __enter__(self):
        _lib.check(_lib.load().pwc_set_half_format(self.fmt), "pwc_set_half_format")

    def __exit__(self, *exc):
        _lib.load().pwc_set_half_format(0)
        return False


def _h16(lib):
    return torch.float16 if lib.pwc_get_half_format() else torch.bfloat16


def warp_corr_lrelu(f1, f2, flow=None, flow_scale=1.0, align_corners=False, mask_thresh=0.9999, slope=0.1,
                    layout="nchw", out=None):
    """Fused warp + mask + 81-channel cost volume + LeakyReLU (pwc_warp_corr_lrelu_fwd).

    layout 'nchw': f1,f2 [B,C,H,W], flow [B,2,H,W] -> out [B,81,H,W] (fp32).
    layout 'nhwc': f1,f2 [B,H,W,C] (last dim may be a wider pitch via .stride(2)), flow [B,H,W,2], out [B,H,W,81]
                   fp32 or bf16 features.
    """
    lib = _lib.load()
    if not f1.is_cuda:
        raise RuntimeError("warp_corr_lrelu: CUDA tensors required (no CPU fallback)")
    st = _lib.stream_handle()
    if layout == "nchw":
        f1, f2 = f1.contiguous(), f2.contiguous()
        B, C, H, W = f1.shape
        flow = None if flow is None else flow.contiguous().float()
        if out is None:
            out = torch.empty(B, 81, H, W, device=f1.device, dtype=torch.float32)
        e = lib.pwc_warp_corr_lrelu_fwd(_lib.ptr(f1), 0, _lib.ptr(f2), 0, _lib.ptr(flow), 0, float(flow_scale),
                                        int(align_corners), mask_thresh, slope, _lib.ptr(out), 0,
                                        B, C, H, W, 4, _lib.NCHW, _lib.F32, st)
    else:
        B, H, W, C = f1.shape
        assert f1.stride(3) == 1 and f2.stride(3) == 1
        dt = _lib.F32 if f1.dtype == torch.float32 else _lib.BF16
        if dt == _lib.BF16 and f1.dtype != _h16(lib):
            raise RuntimeError(f"warp_corr_lrelu: {f1.dtype} features while the library's 16-bit format is {_h16(lib)}")
        flow = None if flow is None else flow.contiguous().float()
        if out is None:
            out = torch.empty(B, H, W, 81, device=f1.device, dtype=f1.dtype)
        e = lib.pwc_warp_corr_lrelu_fwd(_lib.ptr(f1), f1.stride(2), _lib.ptr(f2), f2.stride(2), _lib.ptr(flow), 2,
                                        float(flow_scale), int(align_corners), mask_thresh, slope,
                                        _lib.ptr(out), out.stride(2), B, C, H, W, 4, _lib.NHWC, dt, st)
    _lib.check(e, "pwc_warp_corr_lrelu_fwd")
    return out


def warp(x, flo, flow_scale=1.0, align_corners=False, mask_thresh=0.9999):
    """PWCDCNet.warp (PWCNet.py:141-177): x [B,C,H,W], flo [B,2,H,W] (pixels) -> grid_sample(x, grid + flo) * mask."""
    if not (x.is_cuda and flo.is_cuda):
        raise RuntimeError("warp: CUDA tensors required (no CPU fallback)")
    if x.dim() != 4 or flo.dim() != 4 or flo.shape[1] != 2 or flo.shape[0] != x.shape[0] or flo.shape[2:] != x.shape[2:]:
        raise RuntimeError("warp: x [B,C,H,W] and flo [B,2,H,W] expected")
    if x.dtype != torch.float32 or flo.dtype != torch.float32:
        raise RuntimeError("expected scalar type Float")          # as grid.float() + grid_sample in the reference
    x, flo = x.contiguous(), flo.contiguous()
    B, C, H, W = x.shape
    out = torch.empty_like(x)
    lib = _lib.load()
    with torch.cuda.device_of(x):
        e = lib.pwc_warp_fwd(_lib.ptr(x), _lib.ptr(flo), _lib.ptr(out), B, C, H, W, float(flow_scale), int(bool(align_corners)),
                             float(mask_thresh), _lib.stream_handle())
    _lib.check(e, "pwc_warp_fwd")
    return out


def conv3x3_f32(x, weight, bias, stride=1, dil=1, slope=0.1, in_coff=0, cin_span=None, out=None, out_coff=0,
                chan_map=None, residual=None):
    """fp32 CUDA-core conv (pwc_conv3x3_f32_fwd) on an NHWC tensor x [N,H,W,pitch]; reads channels
    [in_coff, in_coff+cin_span); weight is the torch [Cout,Cin,3,3] tensor (packed here, per call)."""
    lib = _lib.load()
    st = _lib.stream_handle()
    N, H, W, pitch = x.shape
    cout, cin = weight.shape[:2]
    span = cin_span if cin_span is not None else cin
    cinp, coutp = (span + 15) // 16 * 16, (cout + 63) // 64 * 64
    wp = torch.empty(9 * cinp * coutp, device=x.device, dtype=torch.float32)
    mp = None if chan_map is None else torch.tensor(chan_map, device=x.device, dtype=torch.int32)
    w32 = weight.detach().float().contiguous()
    _lib.check(lib.pwc_pack_conv_weight_f32(_lib.ptr(w32), _lib.ptr(wp), _lib.ptr(mp), cout, cin, cinp, coutp, st), "pack")
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    if out is None:
        out = torch.zeros(N, ho, wo, cout, device=x.device, dtype=torch.float32)
    b32 = bias.detach().float().contiguous()
    e = lib.pwc_conv3x3_f32_fwd(_lib.ptr(x), pitch, in_coff, None, 0, 0, span, _lib.ptr(wp), _lib.ptr(b32),
                                _lib.ptr(out), out.shape[-1], out_coff, None, 0, 0, N, cout, N, H, W, stride, dil,
                                slope, _lib.ptr(residual), (residual.shape[-1] if residual is not None else 0), st)
    _lib.check(e, "pwc_conv3x3_f32_fwd")
    torch.cuda.current_stream().synchronize()
    return out


def deconv4x4s2_c2_f32(x, weight, bias, in_coff=0, cin_span=None, chan_map=None):
    """2-channel ConvTranspose2d(4,2,1) on NHWC x [N,H,W,pitch]; weight torch layout [Cin,2,4,4]."""
    lib = _lib.load()
    st = _lib.stream_handle()
    N, H, W, pitch = x.shape
    cin = weight.shape[0]
    span = cin_span if cin_span is not None else (cin + 3) // 4 * 4
    cinp = (span + 3) // 4 * 4
    wp = torch.empty(32 * cinp, device=x.device, dtype=torch.float32)
    mp = None if chan_map is None else torch.tensor(chan_map, device=x.device, dtype=torch.int32)
    w32 = weight.detach().float().contiguous()
    _lib.check(lib.pwc_pack_deconv_weight_f32(_lib.ptr(w32), _lib.ptr(wp), _lib.ptr(mp), cin, cinp, st), "pack")
    out = torch.zeros(N, 2 * H, 2 * W, 2, device=x.device, dtype=torch.float32)
    b32 = bias.detach().float().contiguous()
    e = lib.pwc_deconv4x4s2_c2_f32_fwd(_lib.ptr(x), pitch, in_coff, span, _lib.ptr(wp), _lib.ptr(b32),
                                       _lib.ptr(out), 2, 0, None, 0, N, H, W, st)
    _lib.check(e, "pwc_deconv4x4s2_c2_f32_fwd")
    torch.cuda.current_stream().synchronize()
    return out


def pick_kc(cin_span, stride=1, dil=1, has_in2=False, W=None, coutp=None, out_f32=False):
    """Channel chunk of the tensor-core conv.  The halo kernel (stride 1, dilation <= 16, one input) takes 64-channel
    chunks with a zero-padded partial last chunk; the per-tap kernel needs cin_span % kc == 0."""
    import os
    exact = 64 if cin_span % 64 == 0 else (32 if cin_span % 32 == 0 else 16)
    ydense = int(os.environ.get("PWC_TC_WM_YDENSE_MINDIL", "8")) <= 16 and int(os.environ.get("PWC_TC_WM", "1"))
    # dilation 9..16: the y-dense mode views W as (W/dil, dil) and stores 32-channel groups
    wide = ydense and W is not None and W % dil == 0 and coutp is not None and coutp % 32 == 0 and not out_f32
    halo = int(os.environ.get("PWC_TC_HALO", "1")) and stride == 1 and 1 <= dil <= (16 if wide else 8) and not has_in2
    # 96 channels at dilation 16: three exact 32-channel chunks halve the halo-tile traffic of two padded 64-channel ones
    # (dc_conv5 on B200: 100 us per-tap kernel -> 75 us kc 64 -> 52 us kc 32)
    if halo and dil > 8 and cin_span % 64 and cin_span % 32 == 0 and int(os.environ.get("PWC_DIL16_KC", "32")) == 32:
        return 32
    if halo:
        return 64      # thin layers too: 128-byte rows (SWIZZLE_128B) are what the tensor pipe reads fast
    return exact


def conv3x3_tc(x, weight, bias, stride=1, dil=1, slope=0.1, in_coff=0, cin_span=None, out=None, out_coff=0,
               chan_map=None, out_f32=False, residual=None, reps=1, ts=False, tl=False, head_w=None, partial=None):
    """bf16 tcgen05 conv (pwc_tc_conv_*) on an NHWC bf16 tensor x [N,H,W,pitch]; weight torch [Cout,Cin,3,3] fp32.
    ts=True: tap-stacked mode for 9*Cout <= 128 (fp32 output); partial [N,H,W,>=Cout] fp32 is added before the activation.
    tl=True with head_w [nhead,Cin,3,3] (Cout <= 32): the raw sums of those extra output channels come back as conv3x3_tc.head
    ([N,H,W,nhead] fp32; pwc_tc_conv_set_head_rows)."""
    import ctypes
    lib = _lib.load()
    st = _lib.stream_handle()
    N, H, W, pitch = x.shape
    assert x.dtype == _h16(lib)
    cout, cin = weight.shape[:2]
    span = cin_span if cin_span is not None else (cin + 15) // 16 * 16
    coutp = 128 if (ts or tl) else (cout + 15) // 16 * 16
    kc = pick_kc(span, stride, dil, W=W, coutp=coutp, out_f32=bool(out_f32))
    wp = torch.empty(9 * ((span + kc - 1) // kc) * kc * coutp, device=x.device, dtype=x.dtype)
    mp = None if chan_map is None else torch.tensor(chan_map, device=x.device, dtype=torch.int32)
    w32 = weight.detach().float().contiguous()
    wrows = cout
    if ts:
        assert 9 * cout <= 128 and out_f32 and stride == 1 and dil == 1
        ws = torch.zeros(128, cin, 3, 3, device=x.device, dtype=torch.float32)
        for tap in range(9):
            ws[tap * cout:(tap + 1) * cout, :, 0, 0] = w32[:, :, tap // 3, tap % 3]
        w32, wrows = ws, 128
    if tl:                                      # tap lanes (include/pwc_b200.h, out_f32 | 32): Cout <= 64, 16-bit output
        assert cout <= 64 and not out_f32 and stride == 1 and dil == 1
        ws = torch.zeros(128, cin, 3, 3, device=x.device, dtype=torch.float32)
        if cout <= 32:
            for l in range(3):
                ws[32 * l:32 * l + cout, :, 0, :] = w32[:, :, l, :]
        else:
            ws[0:cout, :, 0, :] = w32[:, :, 0, :]
            ws[64:64 + cout, :, 0, :] = w32[:, :, 1, :]
            ws[0:cout, :, 1, :] = w32[:, :, 2, :]
        if head_w is not None:
            nh = head_w.shape[0]
            assert cout <= 32 and 3 * nh <= 32 and head_w.shape[1] == cin
            for ky in range(3):
                ws[96 + ky * nh:96 + (ky + 1) * nh, :, 0, :] = head_w.detach().float()[:, :, ky, :]
        w32, wrows = ws, 128
    _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(w32), _lib.ptr(wp), _lib.ptr(mp), wrows, cin, span, kc, coutp, st), "pack")
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    if out is None:
        out = torch.zeros(N, ho, wo, ((cout + 31) // 32 * 32 if tl else coutp) if not out_f32 else cout, device=x.device,
                          dtype=torch.float32 if out_f32 else x.dtype)
    b32 = bias.detach().float().contiguous()
    h = ctypes.c_void_p()
    e = lib.pwc_tc_conv_create(ctypes.byref(h), _lib.ptr(x), pitch, in_coff, None, 0, 0, N, span, kc,
                               _lib.ptr(wp), _lib.ptr(b32), cout, coutp, _lib.ptr(out), out.shape[-1], out_coff,
                               None, 0, 0, int(out_f32) | (16 if ts else 0) | (32 if tl else 0), _lib.ptr(residual),
                               (residual.shape[-1] if residual is not None else 0), N, H, W, stride, dil, slope)
    _lib.check(e, "pwc_tc_conv_create")
    try:
        conv3x3_tc.head = None
        if head_w is not None:
            conv3x3_tc.head = torch.full((N, H, W, head_w.shape[0]), 7.0, device=x.device, dtype=torch.float32)
            _lib.check(lib.pwc_tc_conv_set_head_rows(h, _lib.ptr(conv3x3_tc.head), head_w.shape[0]), "pwc_tc_conv_set_head_rows")
        if partial is not None:
            _lib.check(lib.pwc_tc_conv_set_partial(h, _lib.ptr(partial), partial.shape[-1]), "pwc_tc_conv_set_partial")
        _lib.check(lib.pwc_tc_conv_launch(h, st), "pwc_tc_conv_launch")
        if reps > 1:        # timing aid for tools/: CUDA-event time of reps-1 further launches
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps - 1):
                _lib.check(lib.pwc_tc_conv_launch(h, st), "pwc_tc_conv_launch")
            e1.record(); e1.synchronize()
            conv3x3_tc.last_ms = e0.elapsed_time(e1) / (reps - 1)
        torch.cuda.current_stream().synchronize()
    finally:
        lib.pwc_tc_conv_destroy(h)
    return out


def conv3x3_tc_split(x32, weight, bias, stride=1, slope=0.1, split_in=True, split_out=True, reps=1, tl=False):
    """Split-precision tcgen05 conv (pwc_tc_conv_create_split) for tests / tools: x32 [N,H,W,cin] fp32 NHWC is stored as
    hi | lo planes of the current 16-bit format; returns the fp32 reconstruction hi + lo of the output [N,Ho,Wo,cout]."""
    import ctypes
    from . import split as S
    lib = _lib.load()
    st = _lib.stream_handle()
    dt = _h16(lib)
    N, H, W, cin = x32.shape
    cout = weight.shape[0]
    P = (cin + 15) // 16 * 16
    span = P
    coutp = (cout + 31) // 32 * 32 if split_out else (cout + 15) // 16 * 16
    opl = coutp                                  # channels per output plane
    if tl:                                       # tap lanes: 128 MMA rows; output planes of 16 (Cout <= 16) or 32*ceil(Cout/32) channels
        assert cout <= 64 and stride == 1
        opl = 16 if cout <= 16 else (cout + 31) // 32 * 32
        coutp = 128
    kc = pick_kc(span, stride, 1, W=W, coutp=coutp)
    dev = x32.device
    xin = torch.zeros(N, H, W, 2 * P if split_in else P, device=dev, dtype=dt)
    hi, lo = S.planes(x32, dt)
    xin[..., :cin] = hi
    if split_in:
        xin[..., P:P + cin] = lo
    w32 = weight.detach().float().contiguous()
    wrows = cout
    if tl:
        ws = torch.zeros(128, cin, 3, 3, device=dev, dtype=torch.float32)
        if cout <= 32:
            for l in range(3):
                ws[32 * l:32 * l + cout, :, 0, :] = w32[:, :, l, :]
        else:
            ws[0:cout, :, 0, :] = w32[:, :, 0, :]
            ws[64:64 + cout, :, 0, :] = w32[:, :, 1, :]
            ws[0:cout, :, 1, :] = w32[:, :, 2, :]
        w32, wrows = ws, 128
    import os
    fused = split_in and stride == 2 and P == span == kc and kc <= 32 and os.environ.get("PWC_TC_FUSED_SPLIT", "1") != "0"
    kc_pack = kc
    if fused:                                   # adjacent planes in one swizzle row (TcConvParams::fused): [W_hi|W_hi], [W_lo|0]
        wv, cm, vspan, kc_pack = torch.cat([w32, w32, S.residual(w32, dt)], 1).contiguous(), list(range(3 * cin)), 4 * span, 2 * span
    elif split_in:
        wv, cm, vspan = S.virtual_weights(w32.cpu(), dt, kc, span)
        wv = wv.to(dev)
    else:
        wv, cm, vspan = w32, list(range(cin)), (span + kc - 1) // kc * kc
    wp = torch.zeros(9 * vspan * coutp, device=dev, dtype=dt)
    mp = torch.tensor(cm, device=dev, dtype=torch.int32)
    _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(wv), _lib.ptr(wp), _lib.ptr(mp), wrows, wv.shape[1], vspan, kc_pack, coutp, st), "pack")
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    out = torch.zeros(N, ho, wo, 2 * opl if split_out else opl, device=dev, dtype=dt)
    b32 = bias.detach().float().contiguous()
    h = ctypes.c_void_p()
    e = lib.pwc_tc_conv_create_split(ctypes.byref(h), _lib.ptr(xin), xin.shape[-1], 0, None, 0, 0, N, span, kc, _lib.ptr(wp),
                                     _lib.ptr(b32), cout, coutp, _lib.ptr(out), out.shape[-1], 0, None, 0, 0, 32 if tl else 0, None, 0,
                                     N, H, W, stride, 1, slope, P if split_in else 0, opl if split_out else 0, 0, 0)
    _lib.check(e, "pwc_tc_conv_create_split")
    try:
        _lib.check(lib.pwc_tc_conv_launch(h, st), "pwc_tc_conv_launch")
        if reps > 1:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps - 1):
                _lib.check(lib.pwc_tc_conv_launch(h, st), "pwc_tc_conv_launch")
            e1.record(); e1.synchronize()
            conv3x3_tc_split.last_ms = e0.elapsed_time(e1) / (reps - 1)
        torch.cuda.current_stream().synchronize()
    finally:
        lib.pwc_tc_conv_destroy(h)
    res = out[..., :cout].float()
    if split_out:
        res = res + out[..., opl:opl + cout].float()
    return res


def conv3x3_thin_split(x32, weight, bias, slope=0.1, reps=1, stride=1):
    """Split-precision thin layer (pwc_conv3x3_thin_split_fwd): x32 [N,H,W,cin] fp32, cin / cout in {16, 32}; returns hi + lo."""
    from . import split as S
    lib = _lib.load()
    st = _lib.stream_handle()
    dt = _h16(lib)
    N, H, W, cin = x32.shape
    cout = weight.shape[0]
    dev = x32.device
    hi, lo = S.planes(x32, dt)
    xin = torch.cat([hi, lo], -1).contiguous()
    w32 = weight.detach().float()
    wv = torch.cat([w32, S.residual(w32, dt)], 1).contiguous()
    wp = torch.zeros(9 * cout * 2 * cin, device=dev, dtype=dt)
    _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(wv), _lib.ptr(wp), None, cout, 2 * cin, 2 * cin, 2 * cin, cout, st), "pack")
    out = torch.full((N, (H - 1) // stride + 1, (W - 1) // stride + 1, 2 * cout), 7.0, device=dev, dtype=dt)
    b32 = bias.detach().float().contiguous()

    def launch():
        _lib.check(lib.pwc_conv3x3_thin_split_strided_fwd(_lib.ptr(xin), 2 * cin, 0, cin, _lib.ptr(wp), _lib.ptr(b32), cout,
                                                          _lib.ptr(out), 2 * cout, 0, N, H, W, stride, slope, st),
                   "pwc_conv3x3_thin_split_strided_fwd")
    launch()
    if reps > 1:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps - 1):
            launch()
        e1.record(); e1.synchronize()
        conv3x3_thin_split.last_ms = e0.elapsed_time(e1) / (reps - 1)
    torch.cuda.current_stream().synchronize()
    return out[..., :cout].float() + out[..., cout:].float()


def conv1a_bf16(x, weight, bias, slope=0.1):
    """conv1a fused with the im1/im2 split (pwc_conv1a_bf16_fwd): x [B,6,H,W] fp32 -> bf16 NHWC [2B,H/2,W/2,16]."""
    lib = _lib.load()
    B, _, H, W = x.shape
    out = torch.empty(2 * B, H // 2, W // 2, 16, device=x.device, dtype=torch.bfloat16)
    w32, b32 = weight.detach().float().contiguous(), bias.detach().float().contiguous()
    e = lib.pwc_conv1a_bf16_fwd(_lib.ptr(x.contiguous()), _lib.ptr(w32), _lib.ptr(b32), _lib.ptr(out), B, H, W, slope,
                                _lib.stream_handle())
    _lib.check(e, "pwc_conv1a_bf16_fwd")
    torch.cuda.current_stream().synchronize()
    return out


def deconv4x4s2_c2(x, weight, bias, out_dtype=torch.float32, in_coff=0, cin_span=None, chan_map=None):
    """2-channel ConvTranspose2d(4,2,1) with fp32/bf16 input and output (pwc_deconv4x4s2_c2_fwd)."""
    lib = _lib.load()
    st = _lib.stream_handle()
    N, H, W, pitch = x.shape
    cin = weight.shape[0]
    span = cin_span if cin_span is not None else (cin + 3) // 4 * 4
    cinp = (span + 3) // 4 * 4
    wp = torch.empty(32 * cinp, device=x.device, dtype=torch.float32)
    mp = None if chan_map is None else torch.tensor(chan_map, device=x.device, dtype=torch.int32)
    w32 = weight.detach().float().contiguous()
    _lib.check(lib.pwc_pack_deconv_weight_f32(_lib.ptr(w32), _lib.ptr(wp), _lib.ptr(mp), cin, cinp, st), "pack")
    out = torch.zeros(N, 2 * H, 2 * W, 2, device=x.device, dtype=out_dtype)
    outb = torch.zeros(N, 2 * H, 2 * W, 2, device=x.device, dtype=torch.float32)
    b32 = bias.detach().float().contiguous()
    dt = lambda t: _lib.BF16 if t == torch.bfloat16 else _lib.F32  # noqa: E731
    e = lib.pwc_deconv4x4s2_c2_fwd(_lib.ptr(x), dt(x.dtype), pitch, in_coff, span, _lib.ptr(wp), _lib.ptr(b32),
                                   _lib.ptr(out), dt(out_dtype), 2, 0, _lib.ptr(outb), 2, N, H, W, st)
    _lib.check(e, "pwc_deconv4x4s2_c2_fwd")
    torch.cuda.current_stream().synchronize()
    return out, outb


def conv3x3_mma_small(x, weight, bias, stride=1, slope=0.1, in_coff=0, cin_span=None, chan_map=None, x2=None, n_split_in=0,
                      reps=1):
    """Small-level bf16 conv (pwc_conv3x3_mma_fwd) on an NHWC bf16 tensor x [N,H,W,pitch]; weight torch [Cout,Cin,3,3] fp32.
    x2: second input tensor for images >= n_split_in (as the stride-2 pyramid layers read im1 / im2 features)."""
    lib = _lib.load()
    st = _lib.stream_handle()
    if not x.is_cuda:
        raise RuntimeError("conv3x3_mma_small: CUDA tensors required (no CPU fallback)")
    N, H, W, pitch = x.shape
    if x2 is not None:
        N = n_split_in + x2.shape[0]
    assert x.dtype == torch.bfloat16
    cout, cin = weight.shape[:2]
    span = cin_span if cin_span is not None else (cin + 7) // 8 * 8
    coutp = (cout + 15) // 16 * 16
    nchunk = (span + 63) // 64
    wp = torch.empty(9 * nchunk * 64 * coutp, device=x.device, dtype=torch.bfloat16)
    mp = None if chan_map is None else torch.tensor(chan_map, device=x.device, dtype=torch.int32)
    w32 = weight.detach().float().contiguous()
    _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(w32), _lib.ptr(wp), _lib.ptr(mp), cout, cin, span, 64, coutp, st), "pack")
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    out = torch.full((N, ho, wo, coutp), 7.0, device=x.device, dtype=torch.bfloat16)
    b32 = bias.detach().float().contiguous()

    def launch():
        e = lib.pwc_conv3x3_mma_fwd(_lib.ptr(x), pitch, in_coff, _lib.ptr(x2), (x2.shape[-1] if x2 is not None else 0), in_coff,
                                    n_split_in, span, _lib.ptr(wp), _lib.ptr(b32), cout, coutp, _lib.ptr(out), coutp, 0,
                                    None, 0, 0, N, N, H, W, stride, slope, st)
        _lib.check(e, "pwc_conv3x3_mma_fwd")
    launch()
    if reps > 1:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps - 1):
            launch()
        e1.record(); e1.synchronize()
        conv3x3_mma_small.last_ms = e0.elapsed_time(e1) / (reps - 1)
    torch.cuda.current_stream().synchronize()
    return out


def conv3x3_thin(x, weight, bias, stride=1, slope=0.1, in_coff=0, x2=None, n_split_in=0, out=None, out_coff=0, reps=1):
    """Thin-layer bf16 conv (pwc_conv3x3_thin_fwd): x [N,H,W,pitch] bf16 NHWC, weight torch [Cout,Cin,3,3] fp32 with
    Cin, Cout in {16, 32, 64}; x2: second input for images >= n_split_in."""
    lib = _lib.load()
    st = _lib.stream_handle()
    if not x.is_cuda:
        raise RuntimeError("conv3x3_thin: CUDA tensors required (no CPU fallback)")
    N, H, W, pitch = x.shape
    if x2 is not None:
        N = n_split_in + x2.shape[0]
    cout, cin = weight.shape[:2]
    wp = torch.empty(9 * cout * cin, device=x.device, dtype=torch.bfloat16)
    w32 = weight.detach().float().contiguous()
    _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(w32), _lib.ptr(wp), None, cout, cin, cin, cin, cout, st), "pack")
    ho, wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    if out is None:
        out = torch.full((N, ho, wo, cout), 7.0, device=x.device, dtype=torch.bfloat16)
    b32 = bias.detach().float().contiguous()

    def launch():
        e = lib.pwc_conv3x3_thin_fwd(_lib.ptr(x), pitch, in_coff, _lib.ptr(x2), (x2.shape[-1] if x2 is not None else 0), in_coff,
                                     n_split_in, cin, _lib.ptr(wp), _lib.ptr(b32), cout, _lib.ptr(out), out.shape[-1], out_coff,
                                     None, 0, 0, N, N, H, W, stride, slope, st)
        _lib.check(e, "pwc_conv3x3_thin_fwd")
    launch()
    if reps > 1:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps - 1):
            launch()
        e1.record(); e1.synchronize()
        conv3x3_thin.last_ms = e0.elapsed_time(e1) / (reps - 1)
    torch.cuda.current_stream().synchronize()
    return out
