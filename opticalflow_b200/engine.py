"""Host-side execution plan of the B200 PWC-Net forward (PWCDCNet.forward, /root/reference/models/PWCNet.py:180-273).

One ``Plan`` per (batch, H, W): all activation buffers are allocated once (NHWC + pitch, zero-initialised so
padding channels stay zero), weights are repacked once per state_dict version, and the forward is a fixed list of
C-ABI launches (include/pwc_b200.h) on the current CUDA stream -- no torch op, no torch.cat, no allocation.

Concat buffer of level L (replaces the 29 torch.cat of PWCNet.py:202-264), channels:
    [0,32) convL_4 | [32,96) convL_3 | [96,192) convL_2 | [192,320) convL_1 | [320,448) convL_0 |
    [448,529) cost volume | [529,531) up_flow | [531,533) up_feat | [533,544) zero | [544,544+C_L) c1_L
Every layer reads a suffix [start, end) of it and writes the slice in front; all starts are multiples of 32.
The torch input-channel order (corr, c1, up_flow, up_feat) differs from the buffer order only in the tail; the
weight packer applies ``chan_map``.
"""
from __future__ import annotations

import torch

from . import _lib

PYR_CH = [3, 16, 32, 64, 96, 128, 196]          # PWCNet.py:52-69
DEC_OUT = [128, 128, 96, 64, 32]                # PWCNet.py:75-81
DEC_START = [448, 320, 192, 96, 32]             # buffer offset each convL_k READS from
DEC_WRITE = [320, 192, 96, 32, 0]               # buffer offset each convL_k WRITES to
CORR_OFF, UPFLOW_OFF, UPFEAT_OFF, C1_OFF = 448, 529, 531, 544
FLOW_SCALE = {5: 0.625, 4: 1.25, 3: 2.5, 2: 5.0}  # PWCNet.py:212,226,240,256
DC = [("dc_conv1", 128, 1), ("dc_conv2", 128, 2), ("dc_conv3", 128, 4), ("dc_conv4", 96, 8),
      ("dc_conv5", 64, 16), ("dc_conv6", 32, 1)]  # PWCNet.py:126-131


def level_channels(lvl):
    return PYR_CH[lvl]


def concat_width(lvl):
    return C1_OFF if lvl == 6 else C1_OFF + PYR_CH[lvl]


def tail_map(lvl):
    """torch channel order of the decoder input tail [corr | c1 | up_flow | up_feat] (PWCNet.py:215,229,245,259)
    -> offset relative to buffer channel 448."""
    m = list(range(81))
    if lvl < 6:
        C = PYR_CH[lvl]
        m += [C1_OFF - CORR_OFF + i for i in range(C)]
        m += [UPFLOW_OFF - CORR_OFF, UPFLOW_OFF - CORR_OFF + 1, UPFEAT_OFF - CORR_OFF, UPFEAT_OFF - CORR_OFF + 1]
    return m


def chan_map(lvl, start):
    """torch input channel i of a layer reading the concat suffix [start, end) -> channel relative to start."""
    dense = CORR_OFF - start
    return list(range(dense)) + [dense + t for t in tail_map(lvl)]


CONV_MODES = ("fp32", "bf16", "f16", "mixed")


def weights_key(model):
    """Identity of the parameter values a plan has packed.  In-place edits through autograd-visible ops bump
    ``p._version``; re-assigned storage changes ``data_ptr``.  Edits through ``.data`` (``m.weight.data.copy_()``,
    ``nn.init.*(m.weight.data)`` -- the reference's own init idiom, PWCNet.py:134-138) change neither, so the key also
    carries ``model._weights_epoch`` (``PWCDCNet.invalidate_weights()`` bumps it; ``load_state_dict`` does so
    automatically) and, when ``model.verify_weights`` is set, a device-side checksum of every parameter (one reduction
    + one 8-byte D2H copy per forward)."""
    ps = list(model.parameters())
    key = tuple((p.data_ptr(), p._version) for p in ps) + (id(model), getattr(model, "_weights_epoch", 0))
    if getattr(model, "verify_weights", False):
        import torch
        with torch.no_grad():
            sums = torch._foreach_norm([p.detach().double() for p in ps], 1)
            w = torch.arange(1, len(ps) + 1, device=sums[0].device, dtype=torch.float64)
            key += (float((torch.stack(sums) * w).sum().item()),)
    return key


def _roundup(a, b):
    return (a + b - 1) // b * b


class PackedConv:
    __slots__ = ("w", "b", "cin_span", "cout")


class Plan:
    """fp32 execution plan (exact-fp32 mode: every convolution on CUDA cores)."""

    def __init__(self, model, B, H, W, device, align_corners=False):
        if H % 64 or W % 64:
            raise ValueError(f"H and W must be multiples of 64 (6 stride-2 levels), got {H}x{W}")
        self.lib = _lib.load()
        self.B, self.H, self.W, self.device = B, H, W, device
        self.align_corners = int(bool(align_corners))
        z = lambda *s: torch.zeros(*s, device=device, dtype=torch.float32)  # noqa: E731
        self.x4 = z(2 * B, H, W, 4)
        self.hw = {l: (H >> l, W >> l) for l in range(1, 7)}
        self.pa, self.pb = {}, {}
        for l in range(1, 7):
            h, w = self.hw[l]
            self.pa[l] = z(2 * B, h, w, PYR_CH[l])
            self.pb[l] = z(2 * B, h, w, PYR_CH[l])
        self.cat, self.f2, self.flow, self.upflow = {}, {}, {}, {}
        for l in range(2, 7):
            h, w = self.hw[l]
            self.cat[l] = z(B, h, w, concat_width(l))
            self.f2[l] = z(B, h, w, PYR_CH[l])
            self.flow[l] = z(B, h, w, 4)
            if l < 6:
                self.upflow[l] = z(B, h, w, 2)
        self.f1_6 = z(B, *self.hw[6], PYR_CH[6])
        h2, w2 = self.hw[2]
        self.dc = [z(B, h2, w2, c) for (_, c, _) in DC]
        self.flow_out4 = z(B, h2, w2, 4)
        self.out = z(B, 2, h2, w2)
        self.weights_version = None
        self.packed = {}
        self.launches = 0

    # ------------------------------------------------------------------ weights
    def pack_weights(self, model):
        ver = weights_key(model)
        if ver == self.weights_version:
            return
        lib, dev = self.lib, self.device
        st = _lib.stream_handle()
        sd = {k: v.detach().to(device=dev, dtype=torch.float32).contiguous() for k, v in model.state_dict().items()}
        keep = []

        def pack_conv(key, cin_span, cmap):
            wt, bs = sd[key + ".weight"], sd[key + ".bias"]
            cout, cin = wt.shape[0], wt.shape[1]
            cinp, coutp = _roundup(cin_span, 16), _roundup(cout, 64)
            wp = torch.empty(9 * cinp * coutp, device=dev, dtype=torch.float32)
            mp = None
            if cmap is not None:
                assert len(cmap) == cin
                mp = torch.tensor(cmap, device=dev, dtype=torch.int32)
                keep.append(mp)
            _lib.check(lib.pwc_pack_conv_weight_f32(_lib.ptr(wt), _lib.ptr(wp), _lib.ptr(mp), cout, cin, cinp, coutp, st),
                       "pack_conv_weight")
            pc = PackedConv(); pc.w, pc.b, pc.cin_span, pc.cout = wp, bs, cin_span, cout
            self.packed[key] = pc

        def pack_deconv(key, cin_span, cmap):
            wt, bs = sd[key + ".weight"], sd[key + ".bias"]
            cin = wt.shape[0]
            cinp = _roundup(cin_span, 4)
            wp = torch.empty(32 * cinp, device=dev, dtype=torch.float32)
            mp = None
            if cmap is not None:
                mp = torch.tensor(cmap, device=dev, dtype=torch.int32)
                keep.append(mp)
            _lib.check(lib.pwc_pack_deconv_weight_f32(_lib.ptr(wt), _lib.ptr(wp), _lib.ptr(mp), cin, cinp, st),
                       "pack_deconv_weight")
            pc = PackedConv(); pc.w, pc.b, pc.cin_span, pc.cout = wp, bs, cin_span, 2
            self.packed[key] = pc

        for l in range(1, 7):
            names = (f"conv{l}a", f"conv{l}aa", f"conv{l}b") if l < 6 else ("conv6aa", "conv6a", "conv6b")
            pack_conv(names[0] + ".0", 4 if l == 1 else PYR_CH[l - 1], None)
            pack_conv(names[1] + ".0", PYR_CH[l], None)
            pack_conv(names[2] + ".0", PYR_CH[l], None)
        for l in range(2, 7):
            wd = concat_width(l)
            for k in range(5):
                pack_conv(f"conv{l}_{k}.0", wd - DEC_START[k], chan_map(l, DEC_START[k]))
            pack_conv(f"predict_flow{l}", wd, chan_map(l, 0))
            if l > 2:
                pack_deconv(f"deconv{l}", 4, None)
                pack_deconv(f"upfeat{l}", wd, chan_map(l, 0))
        pack_conv("dc_conv1.0", concat_width(2), chan_map(2, 0))
        prev = 128
        for (name, cout, _) in DC[1:]:
            pack_conv(name + ".0", prev, None)
            prev = cout
        pack_conv("dc_conv7", 32, None)
        torch.cuda.current_stream().synchronize()   # chan_map temporaries may be freed after this
        self.weights_version = ver

    # ------------------------------------------------------------------ launches
    def _conv(self, key, src, src_coff, dst, dst_coff, N, hin, win, stride=1, dil=1, slope=0.1,
              src2=None, src2_coff=0, dst2=None, dst2_coff=0, n_split=0, residual=None):
        pc = self.packed[key]
        st = _lib.stream_handle()
        e = self.lib.pwc_conv3x3_f32_fwd(
            _lib.ptr(src), src.shape[-1], src_coff,
            _lib.ptr(src2), (src2.shape[-1] if src2 is not None else 0), src2_coff, pc.cin_span,
            _lib.ptr(pc.w), _lib.ptr(pc.b),
            _lib.ptr(dst), dst.shape[-1], dst_coff,
            _lib.ptr(dst2), (dst2.shape[-1] if dst2 is not None else 0), dst2_coff, n_split,
            pc.cout, N, hin, win, stride, dil, slope,
            _lib.ptr(residual), (residual.shape[-1] if residual is not None else 0), st)
        _lib.check(e, f"conv {key}")
        self.launches += 1

    def _deconv(self, key, src, src_coff, dst, dst_coff, hin, win, dst_b=None):
        pc = self.packed[key]
        e = self.lib.pwc_deconv4x4s2_c2_f32_fwd(
            _lib.ptr(src), src.shape[-1], src_coff, pc.cin_span, _lib.ptr(pc.w), _lib.ptr(pc.b),
            _lib.ptr(dst), dst.shape[-1], dst_coff, _lib.ptr(dst_b), (dst_b.shape[-1] if dst_b is not None else 0),
            self.B, hin, win, _lib.stream_handle())
        _lib.check(e, f"deconv {key}")
        self.launches += 1

    def _corr(self, lvl):
        h, w = self.hw[lvl]
        cat = self.cat[lvl]
        C = PYR_CH[lvl]
        if lvl == 6:
            f1, f1_pitch, flow, fp, scale = self.f1_6, C, None, 0, 1.0
            f1p = _lib.ptr(f1)
        else:
            f1p = _lib.ctypes.c_void_p(cat.data_ptr() + 4 * C1_OFF)
            f1_pitch, flow, fp, scale = cat.shape[-1], self.upflow[lvl], 2, FLOW_SCALE[lvl]
        outp = _lib.ctypes.c_void_p(cat.data_ptr() + 4 * CORR_OFF)
        e = self.lib.pwc_warp_corr_lrelu_fwd(f1p, f1_pitch, _lib.ptr(self.f2[lvl]), C, _lib.ptr(flow), fp, scale,
                                             self.align_corners, 0.9999, 0.1, outp, cat.shape[-1],
                                             self.B, C, h, w, 4, _lib.NHWC, _lib.F32, _lib.stream_handle())
        _lib.check(e, f"warp_corr level {lvl}")
        self.launches += 1

    # ------------------------------------------------------------------ forward
    def run(self, x, all_levels=False):
        """x: [B,6,H,W] fp32 CUDA contiguous.  Returns flow2 [B,2,H/4,W/4] (a plan-owned tensor)."""
        B, lib = self.B, self.lib
        self.launches = 0
        _lib.check(lib.pwc_pack_input_f32(_lib.ptr(x), _lib.ptr(self.x4), B, self.H, self.W, _lib.stream_handle()), "pack_input")
        self.launches += 1
        src, (hin, win) = self.x4, (self.H, self.W)
        for l in range(1, 7):
            names = (f"conv{l}a", f"conv{l}aa", f"conv{l}b") if l < 6 else ("conv6aa", "conv6a", "conv6b")
            h, w = self.hw[l]
            if l <= 2:      # previous level's output lives in one 2B-image buffer
                self._conv(names[0] + ".0", src, 0, self.pa[l], 0, 2 * B, hin, win, stride=2)
            else:           # image-1 features live in the concat buffer, image-2 features in f2
                self._conv(names[0] + ".0", self.cat[l - 1], C1_OFF, self.pa[l], 0, 2 * B, hin, win, stride=2,
                           src2=self.f2[l - 1], n_split=B)
            self._conv(names[1] + ".0", self.pa[l], 0, self.pb[l], 0, 2 * B, h, w)
            if l == 1:
                self._conv(names[2] + ".0", self.pb[l], 0, self.pa[l], 0, 2 * B, h, w)
                src = self.pa[l]
            elif l < 6:
                self._conv(names[2] + ".0", self.pb[l], 0, self.cat[l], C1_OFF, 2 * B, h, w, dst2=self.f2[l], n_split=B)
            else:
                self._conv(names[2] + ".0", self.pb[l], 0, self.f1_6, 0, 2 * B, h, w, dst2=self.f2[l], n_split=B)
            hin, win = h, w
        for l in (6, 5, 4, 3, 2):
            h, w = self.hw[l]
            cat = self.cat[l]
            self._corr(l)
            for k in range(5):
                self._conv(f"conv{l}_{k}.0", cat, DEC_START[k], cat, DEC_WRITE[k], B, h, w)
            self._conv(f"predict_flow{l}", cat, 0, self.flow[l], 0, B, h, w, slope=1.0)
            if l > 2:
                self._deconv(f"deconv{l}", self.flow[l], 0, self.cat[l - 1], UPFLOW_OFF, h, w, dst_b=self.upflow[l - 1])
                self._deconv(f"upfeat{l}", cat, 0, self.cat[l - 1], UPFEAT_OFF, h, w)
        h, w = self.hw[2]
        src, coff = self.cat[2], 0
        for i, (name, cout, dil) in enumerate(DC):
            self._conv(name + ".0", src, 0, self.dc[i], 0, B, h, w, dil=dil)
            src = self.dc[i]
        self._conv("dc_conv7", src, 0, self.flow_out4, 0, B, h, w, slope=1.0, residual=self.flow[2])
        _lib.check(lib.pwc_unpack_flow_f32(_lib.ptr(self.flow_out4), 4, 0, _lib.ptr(self.out), B, h, w, 1.0,
                                           _lib.stream_handle()), "unpack_flow")
        self.launches += 1
        return self.out

    def level_flow(self, lvl):
        """flowL as NCHW (training-mode tuple of PWCNet.py:270-271 and tests)."""
        return self.flow[lvl][..., :2].permute(0, 3, 1, 2).contiguous()


def make_plan(model, B, H, W, device, conv_mode="fp32", align_corners=False, video=False):
    if conv_mode == "fp32":
        if video:
            raise ValueError("the shared-pyramid video plan exists for the 16-bit conv modes only")
        return Plan(model, B, H, W, device, align_corners)
    if conv_mode in ("bf16", "f16"):
        from .engine_bf16 import PlanBF16
        return PlanBF16(model, B, H, W, device, align_corners, video=video, half=conv_mode)
    if conv_mode == "mixed":
        from .engine_mixed import PlanMixed
        return PlanMixed(model, B, H, W, device, align_corners, video=video)
    raise ValueError(f"unknown conv_mode {conv_mode!r} (expected one of {CONV_MODES})")
