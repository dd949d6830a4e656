"""pwc_tc_conv_launch_static_splitk: the split-K tcgen05 conv launch that allocates nothing (scratch in a file-scope
__device__ array; include/pwc_b200.h).  Checked through the C ABI against the existing entry point
pwc_tc_conv_launch on the same handle, two calls in a row, at the tolerance of the other tcgen05 conv tests
(bf16 output rounding, 1e-2 of the tensor scale, tests/test_gpu_bf16.py::test_conv3x3_tc_matches_torch).
Collected last on purpose."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _make(cin, cout, N, H, W, dil, seed):
    from opticalflow_b200 import _lib, ops
    lib = _lib.load()
    d = torch.device("cuda:0")
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, cin, H, W, generator=g).bfloat16()
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).bfloat16().float()
    b = torch.randn(cout, generator=g) * 0.1
    ref = F.leaky_relu(F.conv2d(x.float(), w, b, padding=dil, dilation=dil), 0.1)
    span = (cin + 15) // 16 * 16
    coutp = (cout + 15) // 16 * 16
    xn = torch.zeros(N, H, W, span, dtype=torch.bfloat16)
    xn[..., :cin] = x.permute(0, 2, 3, 1)
    xn = xn.to(d)
    kc = ops.pick_kc(span, 1, dil, W=W, coutp=coutp, out_f32=False)
    wp = torch.zeros(9 * ((span + kc - 1) // kc) * kc * coutp, device=d, dtype=torch.bfloat16)
    st = _lib.stream_handle()
    wd, bd = w.to(d).contiguous(), b.to(d).contiguous()
    _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(wd), _lib.ptr(wp), None, cout, cin, span, kc, coutp, st), "pack")
    out = torch.zeros(N, H, W, coutp, device=d, dtype=torch.bfloat16)
    h = ctypes.c_void_p()
    e = lib.pwc_tc_conv_create(ctypes.byref(h), _lib.ptr(xn), span, 0, None, 0, 0, N, span, kc, _lib.ptr(wp), _lib.ptr(bd),
                               cout, coutp, _lib.ptr(out), coutp, 0, None, 0, 0, 0, None, 0, N, H, W, 1, dil, 0.1)
    _lib.check(e, "pwc_tc_conv_create")
    keep = (xn, wp, wd, bd)
    return lib, h, out, ref, keep


@pytest.mark.parametrize("case", [
    # cin, cout, N, H, W, dil, ksplit     level-5 / level-6 decoder shapes (few pixel tiles, long K loops)
    (256, 128, 2, 14, 32, 1, 0), (565, 96, 1, 7, 16, 1, 4), (213, 64, 2, 9, 21, 1, 2), (128, 128, 1, 14, 32, 2, 0),
])
def test_static_splitk_matches_plain_launch(case):
    from opticalflow_b200 import _lib
    cin, cout, N, H, W, dil, ksplit = case
    lib, h, out, ref, keep = _make(cin, cout, N, H, W, dil, 900 + cin + cout)
    try:
        st = _lib.stream_handle()
        _lib.check(lib.pwc_tc_conv_launch(h, st), "pwc_tc_conv_launch")
        torch.cuda.synchronize()
        plain = out.clone()
        scale = ref.abs().max().item()
        got = plain.float().permute(0, 3, 1, 2).cpu()[:, :cout]
        assert (got - ref).abs().max().item() <= 1e-2 * scale          # the existing entry point itself is right
        for call in range(2):                                          # two calls in a row share the static array
            out.fill_(7.0)
            _lib.check(lib.pwc_tc_conv_launch_static_splitk(h, ksplit, st), "pwc_tc_conv_launch_static_splitk")
            torch.cuda.synchronize()
            diff = (out.float() - plain.float()).abs().max().item()
            assert diff <= 1e-2 * scale, (call, diff, scale)
            got = out.float().permute(0, 3, 1, 2).cpu()[:, :cout]
            assert (got - ref).abs().max().item() <= 1e-2 * scale, call
            if out.shape[-1] > cout:
                assert out[..., cout:].abs().max().item() == 0.0       # pad channels are written as zeros
    finally:
        lib.pwc_tc_conv_destroy(h)


def test_static_splitk_rejects_what_it_cannot_hold():
    from opticalflow_b200 import _lib
    # 8 x 112 x 256 pixels x 128 channels of fp32 partial sums = 117 MB > the 16 MiB static array
    lib, h, out, ref, keep = _make(128, 128, 8, 112, 256, 1, 5)
    try:
        e = lib.pwc_tc_conv_launch_static_splitk(h, 2, _lib.stream_handle())
        assert e == 1                                                   # cudaErrorInvalidValue, nothing launched
        torch.cuda.synchronize()
        assert out.abs().max().item() == 0.0
    finally:
        lib.pwc_tc_conv_destroy(h)
    # single-chunk layer: no K range to split
    lib, h, out, ref, keep = _make(64, 64, 1, 14, 32, 1, 6)
    try:
        assert lib.pwc_tc_conv_launch_static_splitk(h, 2, _lib.stream_handle()) == 1
        assert lib.pwc_tc_conv_launch_static_splitk(None, 2, _lib.stream_handle()) == 1
    finally:
        lib.pwc_tc_conv_destroy(h)
