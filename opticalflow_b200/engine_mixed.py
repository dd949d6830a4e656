"""conv_mode 'mixed': the 16-bit tensor-core plan with fp32-grade precision where the flow error is made
(PWCDCNet.forward, /root/reference/models/PWCNet.py:180-273; error budget: tools/error_budget*.py).

Measured on the CPU oracle with per-tensor roundings (1024x448, reference init): an all-half forward is 0.04-0.05 px (mean
EPE delta on the x20 flow) from the fp32 reference; 80% of that comes from the feature pyramid (every rounding there is
amplified by five coarse-to-fine warps), most of the rest from the level 4-6 decoders and from the up_flow / up_feat inputs
of each decoder.  The cost volume itself tolerates half inputs (0.0005 px).  So this plan keeps, in SPLIT precision (two
half planes hi + lo per tensor, three tensor-core products per K step: include/pwc_b200.h "Split precision"):
  * the whole feature pyramid (conv1a .. conv6b) -- the cost volumes and the c1 decoder inputs read its hi plane;
  * the level 5 and 4 decoders (conv{5,4}_0..4), their flow / upfeat heads and concat buffers (measured on the B200, mean
    EPE delta at 1024x448 for three inputs: no split decoder level 0.0196 / 0.0133 / 0.0194 px, level 4 only 0.0160 / 0.0110 /
    0.0159, levels 5+4 0.0131 / 0.0091 / 0.0130, levels 6+5+4 0.0130 / 0.0090 / 0.0127: level 6 buys nothing);
  * up_flow / up_feat of levels 3 and 2, as three 16-bit copies (hi, lo, hi) in the spare channels of the concat buffer
    ("K padding": the single-precision consumers read them with weights W_hi, W_hi, W_lo);
and runs the level 3 / 2 decoders and the context network -- 80% of the FLOPs -- as the 'f16' plan does.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from . import split as S
from .engine import (C1_OFF, CORR_OFF, DC, DEC_START, DEC_WRITE, FLOW_SCALE, PYR_CH, UPFEAT_OFF, UPFLOW_OFF, chan_map,
                     concat_width)
from .engine_bf16 import (PlanBF16, _TcLayer, _pfup_weights, _roundup, _s2d_weights, S2D_DEC, S2D_DEC_MAXCOUT, S2D_DEC_MAXLVL,
                          TL_MODE, TL_MAXLVL)

CPM = [4, 16, 32, 64, 96, 128, 224]     # channels per PLANE of the pyramid tensors (level 6: 196 -> 128 + 96 output groups)
import os
# decoder levels kept in split precision (PWC_MIXED_LEVELS: error-budget experiments, see the module docstring)
SPLIT_LEVELS = tuple(int(v) for v in os.environ.get("PWC_MIXED_LEVELS", "5,4").split(",") if v)
# the split decoder levels keep their WEIGHTS in single precision (two products per K step instead of three): measured on
# the B200 0.0131 -> 0.0149 / 0.0091 -> 0.0103 / 0.0130 -> 0.0151 px at 1024x448 for a third less tensor work there; dropping
# W_lo in the pyramid instead costs 0.025-0.028 px (PWC_MIXED_DEC_PRODUCTS=3 restores the full split)
DEC_PRODUCTS = int(os.environ.get("PWC_MIXED_DEC_PRODUCTS", "2"))
# levels 1-2 stride-1 layers on the tcgen05 tap-lane kernel instead of the mma.sync one: measured 458 / 120 us against 89 / 77 us.
# Not a weight-streaming problem (a narrow resident-weight chunk gave 450 us): the ncu source page (tools/prof_tl.py) shows the
# three epilogue warps issue ~3500 instructions per 240-pixel tile at ~0.3 IPC each (96 fp32 TMEM values summed through shared
# memory per 16-channel output pixel, then the hi/lo split), 16.6k cycles per tile against 2.6k of MMA issue -- off
THIN_TC = os.environ.get("PWC_MIXED_THIN_TC", "0") != "0"
THIN_S2 = os.environ.get("PWC_MIXED_THIN_S2", "1") != "0"
# the flow head's sums over the 533 concat channels conv{l}_4 reads ride along in that layer's idle MMA rows.  Level 2 only by default
# (conv2_4 90 -> 94 us, predict_flow2 71 -> 22 us); at level 3 the 10-channel flow + upfeat head costs conv3_4 9 us and saves pfup3 8
# (that head is bound by its scattered 2-byte upfeat stores, not by its input): PWC_MIXED_HEAD_ROWS_LEVELS=2,3 turns it on there too
HEAD_ROWS = os.environ.get("PWC_MIXED_HEAD_ROWS", "1") != "0"
HEAD_ROWS_LEVELS = tuple(int(v) for v in os.environ.get("PWC_MIXED_HEAD_ROWS_LEVELS", "2").split(",") if v)
KP_LO, KP_DUP = 4, 8                    # K padding: lo copies of up_flow|up_feat at +4 (533..536), second hi copies at +8


def _kpad(base, lvl, start):
    """weights of a single-precision layer reading the concat suffix [start, end) of level `lvl` (3 or 2), extended by the
    lo / second-hi copies of its 4 up_flow | up_feat input channels.  base: key or callable -> (w, b)."""
    cm = chan_map(lvl, start)
    n = len(cm)
    idx = list(range(n - 4, n))                                   # torch order: [corr | c1 | up_flow(2) | up_feat(2)]
    lo_pos = [UPFLOW_OFF - start + KP_LO + j for j in range(4)]
    dup_pos = [UPFLOW_OFF - start + KP_DUP + j for j in range(4)]

    def make(sd, dtype):
        w, b = base(sd) if callable(base) else (sd[base + ".weight"], sd[base + ".bias"])
        w_ext, _ = S.kpad_weights(w, dtype, idx, lo_pos, dup_pos, cm)
        return w_ext, b
    return make, cm + lo_pos + dup_pos


class PlanMixed(PlanBF16):
    def __init__(self, model, B, H, W, device, align_corners=False, use_graph=True, video=False):
        super().__init__(model, B, H, W, device, align_corners, use_graph, video, half="f16")
        NP = self.NP
        zh = lambda *s: torch.zeros(*s, device=device, dtype=self.hdt)      # noqa: E731
        # pyramid ping-pong tensors, [hi | lo] planes per pixel
        self.pa = {l: zh(NP, *self.hw[l], 2 * CPM[l]) for l in range(1, 7)}
        self.pb = {l: zh(NP, *self.hw[l], 2 * CPM[l]) for l in range(1, 7)}
        self.f2 = {l: None for l in range(2, 7)}
        self.f1_6 = None
        for l in SPLIT_LEVELS:                                              # split concat buffers: [hi plane | lo plane]
            self.cat[l] = zh(B, *self.hw[l], 2 * concat_width(l))

    # ------------------------------------------------------------------ layers without a tcgen05 handle
    def _plain(self, key, cin, cout, wshape, wfn, launch, flops):
        L = _TcLayer()
        L.kkskip, L.groups, L.handle, L.cmap, L.split_in, L.kpad = False, None, None, None, False, None
        L.key, L.cin, L.cout, L.wfn = key, cin, cout, wfn
        L.w = torch.zeros(wshape, device=self.device, dtype=self.hdt)
        L.b = torch.zeros(cout, device=self.device, dtype=torch.float32)
        self.layers[key] = L
        self.op_info.append((key, flops))
        self.ops.append(launch(L))
        return L

    def _thin_split(self, key, C, src, dst, N, h, w, cout=None, stride=1):
        """16 -> 16 / 32 -> 32 stride-1 (or 16 -> 32 stride-2: conv2a) pyramid layer on [hi | lo] tensors
        (pwc_conv3x3_thin_split_strided_fwd); h, w: input dims."""
        lib = self.lib
        cout = C if cout is None else cout

        def wfn(sd):
            wt = sd[key + ".weight"]
            return torch.cat([wt, S.residual(wt, self.hdt)], 1).contiguous(), sd[key + ".bias"]

        def launch(L):
            args = (_lib.ptr(src), src.shape[-1], 0, C, _lib.ptr(L.w), _lib.ptr(L.b), cout, _lib.ptr(dst), dst.shape[-1], 0, N, h, w,
                    stride, 0.1)
            return lambda st: _lib.check(lib.pwc_conv3x3_thin_split_strided_fwd(*args, st), key)
        ho, wo = (h - 1) // stride + 1, (w - 1) // stride + 1
        L = self._plain(key, 2 * C, cout, 9 * cout * 2 * C, wfn, launch, 2.0 * N * ho * wo * C * cout * 9)
        L.span, L.kc, L.coutp, L.wrows = 2 * C, 2 * C, cout, cout

    def _copy(self, name, src, src_coff, dst, dst_coff, C, n_img, h, w):
        lib = self.lib
        args = (_lib.ptr(src), src.shape[-1], src_coff, _lib.ptr(dst), dst.shape[-1], dst_coff, C, n_img * h * w)
        self.op_info.append((name, 0.0))
        self.ops.append(lambda st: _lib.check(lib.pwc_copy_channels_h16(*args, st), name))

    def _deconv_split(self, key, src, dst, dst_coff, hin, win, lo_off, dup_off, dst_b):
        dev, lib, B = self.device, self.lib, self.B
        w = torch.zeros(32 * 4, device=dev, dtype=torch.float32)
        b = torch.zeros(2, device=dev, dtype=torch.float32)
        self.small[key] = (w, b, 2, 4)
        args = (_lib.ptr(src), _lib.F32, src.shape[-1], 0, 4, _lib.ptr(w), _lib.ptr(b), _lib.ptr(dst), _lib.BF16, dst.shape[-1],
                dst_coff, _lib.ptr(dst_b), dst_b.shape[-1], B, hin, win, lo_off, dup_off)
        self.op_info.append((key, 2.0 * B * hin * win * 4 * 2 * 2 * 4))
        self.ops.append(lambda st: _lib.check(lib.pwc_deconv4x4s2_c2_split_fwd(*args, st), key))

    def _corr_mixed(self, lvl):
        """fused warp + cost volume on the HI planes of the pyramid features (half inputs cost 0.0005 px), half output"""
        h, w = self.hw[lvl]
        cat, C, B = self.cat[lvl], PYR_CH[lvl], self.B
        feat = self.pa[lvl]                                   # 'b' layer output: images [0,B) = image 1, [B,2B) = image 2
        pitch = feat.shape[-1]
        img = h * w * pitch * 2
        f1p = ctypes.c_void_p(feat.data_ptr())
        f2p = ctypes.c_void_p(feat.data_ptr() + (img if self.video else B * img))
        flow, fp, scale = (None, 0, 1.0) if lvl == 6 else (self.upflow[lvl], 2, FLOW_SCALE[lvl])
        outp = ctypes.c_void_p(cat.data_ptr() + 2 * CORR_OFF)
        lib = self.lib
        args = (f1p, pitch, f2p, pitch, _lib.ptr(flow), fp, scale, self.align_corners, 0.9999, 0.1, outp, cat.shape[-1],
                B, C, h, w, 4, _lib.NHWC, _lib.BF16)
        self.op_info.append((f"warp_corr{lvl}", 2.0 * B * h * w * 81 * C))
        self.ops.append(lambda st: _lib.check(lib.pwc_warp_corr_lrelu_fwd(*args, st), f"warp_corr{lvl}"))

    # ------------------------------------------------------------------ the launch list
    def _build(self):
        B, lib, dev = self.B, self.lib, self.device
        NP, vid = self.NP, self.video
        self.w1a = torch.zeros(16, 3, 3, 3, device=dev)
        self.b1a = torch.zeros(16, device=dev)
        a1 = (_lib.ptr(self.x_static), _lib.ptr(self.w1a), _lib.ptr(self.b1a), _lib.ptr(self.pa[1]), (NP if vid else B), self.H,
              self.W, 0.1, int(vid))
        self.op_info.append(("conv1a.0", 2.0 * NP * (self.H // 2) * (self.W // 2) * 16 * 27))
        self.ops.append(lambda st: _lib.check(lib.pwc_conv1a_split_fwd(*a1, st), "conv1a"))
        self._conv1a_on = lambda x, st: _lib.check(lib.pwc_conv1a_split_fwd(_lib.ptr(x), *a1[1:], st), "conv1a")
        # ---- feature pyramid, split precision: a: pa[l-1] -> pa[l] (stride 2), aa: pa[l] -> pb[l], b: pb[l] -> pa[l]
        for l in range(1, 7):
            names = (f"conv{l}a", f"conv{l}aa", f"conv{l}b") if l < 6 else ("conv6aa", "conv6a", "conv6b")
            h, w = self.hw[l]
            C, CP = PYR_CH[l], CPM[l]
            if l == 2 and THIN_S2:
                # conv2a (16 -> 32, stride 2) on the mma.sync thin kernel: the tcgen05 per-tap kernel's stride-2 gather is bound by
                # the TMA's ~6 cycles per 64-byte pixel row (88 us)
                hin, win = self.hw[1]
                self._thin_split(names[0] + ".0", PYR_CH[1], self.pa[1], self.pa[2], NP, hin, win, cout=C, stride=2)
            elif l > 1:
                hin, win = self.hw[l - 1]
                self._tc(names[0] + ".0", PYR_CH[l - 1], C, self.pa[l - 1], 0, PYR_CH[l - 1], self.pa[l], 0, NP, hin, win, stride=2,
                         in_lo=CPM[l - 1], out_lo=CP, no_small=True)
            for name, src, dst in ((names[1], self.pa[l], self.pb[l]), (names[2], self.pb[l], self.pa[l])):
                key = name + ".0"
                if C <= 32 and THIN_TC:
                    # (experiment, off by default) three tap lanes, split in and out; the mma.sync form below runs 3 x 8.5 GFLOP at
                    # 290-330 TFLOP/s (54% of the legacy tensor pipe) = 77-89 us per layer
                    self._tc(key, C, C, src, 0, C, dst, 0, NP, h, w, in_lo=CP, out_lo=CP, no_small=True, tl=True)
                elif C <= 32:
                    self._thin_split(key, C, src, dst, NP, h, w)
                elif C <= 128:
                    self._tc(key, C, C, src, 0, C, dst, 0, NP, h, w, in_lo=CP, out_lo=CP, no_small=True)
                else:                                       # 196 outputs: two launches of <= 128 output channels
                    fl = 2.0 * NP * h * w * C * C * 9
                    for part, (r0, r1) in enumerate(((0, 128), (128, C))):
                        sl = (lambda k, a, b: (lambda sd: (sd[k + ".weight"][a:b], sd[k + ".bias"][a:b])))(key, r0, r1)
                        self._tc(f"{key}#{part}", C, r1 - r0, src, 0, C, dst, r0, NP, h, w, in_lo=CP, out_lo=CP, wfn=sl,
                                 flops=fl * (r1 - r0) / C, no_small=True)
        # ---- decoders
        self.head_part = {}
        for l in (6, 5, 4, 3, 2):
            h, w = self.hw[l]
            cat, wd = self.cat[l], concat_width(l)
            sp = l in SPLIT_LEVELS
            lo = wd if sp else 0
            if l < 6:                                       # c1 (image-1 features) into its concat slot: hi plane (+ lo if split)
                self._copy(f"c1_copy{l}", self.pa[l], 0, cat, C1_OFF, PYR_CH[l], B, h, w)
                if sp:
                    self._copy(f"c1_copy{l}_lo", self.pa[l], CPM[l], cat, wd + C1_OFF, PYR_CH[l], B, h, w)
            self._corr_mixed(l)
            for k in range(5):
                cm = chan_map(l, DEC_START[k])
                cout = [128, 128, 96, 64, 32][k]
                key = f"conv{l}_{k}.0"
                if sp:
                    self._tc(key, len(cm), cout, cat, DEC_START[k], wd - DEC_START[k], cat, DEC_WRITE[k], B, h, w, cmap=cm,
                             in_lo=(-lo if DEC_PRODUCTS == 2 else lo), out_lo=lo, no_small=True)
                    continue
                if l == 6:                                  # no up_flow / up_feat inputs at the coarsest level: the f16 plan's layer
                    self._tc(key, len(cm), cout, cat, DEC_START[k], wd - DEC_START[k], cat, DEC_WRITE[k], B, h, w, cmap=cm)
                    continue
                kp, cme = _kpad(key, l, DEC_START[k])
                kpf = (lambda f: (lambda sd: f(sd, self.hdt)))(kp)
                g = 128 // cout
                if TL_MODE and cout <= 64 and l <= TL_MAXLVL:
                    head = None
                    if k == 4 and HEAD_ROWS and l in HEAD_ROWS_LEVELS and DEC_START[4] == 32:
                        # the level's head (predict_flow2 | predict_flow + upfeat) over this layer's 533 input channels
                        nh = 2 if l == 2 else 10
                        hb = f"predict_flow{l}" if l == 2 else _pfup_weights(l)
                        hsuf = (lambda b_: (lambda sd: (lambda wb: (wb[0][:, 32:].contiguous(), wb[1]))(
                            b_(sd) if callable(b_) else (sd[b_ + ".weight"], sd[b_ + ".bias"]))))(hb)
                        hk, hcm = _kpad(hsuf, l, DEC_START[4])
                        assert hcm == cme
                        self.head_part[l] = torch.zeros(B, h, w, nh, device=self.device, dtype=torch.float32)
                        head = ((lambda f: (lambda sd: f(sd, self.hdt)))(hk), nh, self.head_part[l])
                    self._tc(key, len(cme), cout, cat, DEC_START[k], wd - DEC_START[k], cat, DEC_WRITE[k], B, h, w, cmap=cme, wfn=kpf,
                             flops=2.0 * B * h * w * cout * len(cm) * 9, tl=True, head=head)
                    continue
                if S2D_DEC and cout <= S2D_DEC_MAXCOUT and l <= S2D_DEC_MAXLVL and w % g == 0:
                    v = cat[:B].view(B, h, w // g, g * wd)
                    cm2 = [dxp * wd + DEC_START[k] + c for dxp in range(g) for c in cme]
                    self._tc(key, g * len(cme), g * cout, v, 0, g * wd, v, DEC_WRITE[k], B, h, w // g, cmap=cm2,
                             wfn=_s2d_weights(kpf, g), flops=2.0 * B * h * w * cout * len(cm) * 9, groups=(g, wd))
                    continue
                self._tc(key, len(cme), cout, cat, DEC_START[k], wd - DEC_START[k], cat, DEC_WRITE[k], B, h, w, cmap=cme, wfn=kpf,
                         flops=2.0 * B * h * w * cout * len(cm) * 9)
            cm = chan_map(l, 0)
            if l > 2:
                tgt = l - 1
                tsp = tgt in SPLIT_LEVELS
                twd = concat_width(tgt)
                up_lo, up_dup = (twd, 0) if tsp else (KP_LO, KP_DUP)
                if sp:
                    self._tc(f"pfup{l}", len(cm), 10, cat, 0, wd, self.flow[l], 0, B, h, w, slope=1.0, cmap=cm, out_f32=2,
                             dst2=self.cat[tgt], dst2_coff=UPFEAT_OFF, wfn=_pfup_weights(l), ts=True,
                             in_lo=lo, up_lo=up_lo, up_dup=up_dup)     # the flow / upfeat heads keep all three products
                elif l == 6:
                    self._tc(f"pfup{l}", len(cm), 10, cat, 0, wd, self.flow[l], 0, B, h, w, slope=1.0, cmap=cm, out_f32=2,
                             dst2=self.cat[tgt], dst2_coff=UPFEAT_OFF, wfn=_pfup_weights(l), ts=True, up_lo=up_lo, up_dup=up_dup)
                elif l in self.head_part:
                    first32 = (lambda f: (lambda sd: (lambda wb: (wb[0][:, :32].contiguous(), wb[1]))(f(sd))))(_pfup_weights(l))
                    self._tc(f"pfup{l}", 32, 10, cat, 0, 32, self.flow[l], 0, B, h, w, slope=1.0, out_f32=2,
                             dst2=self.cat[tgt], dst2_coff=UPFEAT_OFF, wfn=first32, ts=True, up_lo=up_lo, up_dup=up_dup,
                             flops=2.0 * B * h * w * 10 * len(cm) * 9, partial=self.head_part[l])
                else:
                    kp, cme = _kpad(_pfup_weights(l), l, 0)
                    self._tc(f"pfup{l}", len(cme), 10, cat, 0, wd, self.flow[l], 0, B, h, w, slope=1.0, cmap=cme, out_f32=2,
                             dst2=self.cat[tgt], dst2_coff=UPFEAT_OFF, wfn=(lambda f: (lambda sd: f(sd, self.hdt)))(kp), ts=True,
                             up_lo=up_lo, up_dup=up_dup, flops=2.0 * B * h * w * 10 * len(cm) * 9)
                self._deconv_split(f"deconv{l}", self.flow[l], self.cat[tgt], UPFLOW_OFF, h, w, up_lo, up_dup, self.upflow[tgt])
            elif l in self.head_part:
                first32 = (lambda k_: (lambda sd: (sd[k_ + ".weight"][:, :32].contiguous(), sd[k_ + ".bias"])))(f"predict_flow{l}")
                self._tc(f"predict_flow{l}", 32, 2, cat, 0, 32, self.flow[l], 0, B, h, w, slope=1.0, out_f32=True, ts=True,
                         wfn=first32, flops=2.0 * B * h * w * 2 * len(cm) * 9, partial=self.head_part[l])
            else:
                kp, cme = _kpad(f"predict_flow{l}", l, 0)
                self._tc(f"predict_flow{l}", len(cme), 2, cat, 0, wd, self.flow[l], 0, B, h, w, slope=1.0, cmap=cme, out_f32=True,
                         ts=True, wfn=(lambda f: (lambda sd: f(sd, self.hdt)))(kp), flops=2.0 * B * h * w * 2 * len(cm) * 9)
        # ---- context network (single precision, as the f16 plan) ----
        h, w = self.hw[2]
        kp, cme = _kpad("dc_conv1.0", 2, 0)
        src, cin, span, cmap, wfn = self.cat[2], len(cme), concat_width(2), cme, (lambda f: (lambda sd: f(sd, self.hdt)))(kp)
        for i, (name, cout, dil) in enumerate(DC):
            g = 128 // cout if cout <= S2D_DEC_MAXCOUT else 1
            if TL_MODE and cout <= 64 and dil == 1 and span >= 128:
                self._tc(name + ".0", cin, cout, src, 0, span, self.dc[i], 0, B, h, w, cmap=cmap, wfn=wfn, tl=True)
            elif S2D_DEC and g > 1 and dil == 1 and cmap is None and w % g == 0 and src.shape[-1] == span:
                vi, vo = src[:B].view(B, h, w // g, g * span), self.dc[i].view(B, h, w // g, g * cout)
                self._tc(name + ".0", g * cin, g * cout, vi, 0, g * span, vo, 0, B, h, w // g,
                         wfn=_s2d_weights(name + ".0", g), flops=2.0 * B * h * w * cout * cin * 9, groups=(g, cout))
            else:
                self._tc(name + ".0", cin, cout, src, 0, span, self.dc[i], 0, B, h, w, dil=dil, cmap=cmap, wfn=wfn,
                         flops=2.0 * B * h * w * cout * (cin - (8 if i == 0 else 0)) * 9)
            src, cin, span, cmap, wfn = self.dc[i], cout, cout, None, None
        L = _TcLayer()
        L.kkskip, L.groups, L.handle, L.cmap, L.split_in, L.kpad = False, None, None, None, False, None
        L.key, L.cin, L.cout, L.span, L.coutp, L.kc, L.wrows = "dc_conv7", 32, 2, 32, 16, 32, 16

        def pad16(sd):
            wt = torch.zeros(16, 32, 3, 3, device=dev, dtype=torch.float32)
            wt[:2] = sd["dc_conv7.weight"]
            return wt, sd["dc_conv7.bias"]
        L.wfn = pad16
        L.w = torch.zeros(9 * 32 * 16, device=dev, dtype=self.hdt)
        L.b = torch.zeros(2, device=dev, dtype=torch.float32)
        self.layers["dc_conv7"] = L
        ha = (_lib.ptr(src), src.shape[-1], 0, 32, _lib.ptr(L.w), _lib.ptr(L.b), _lib.ptr(self.flow[2]), self.flow[2].shape[-1],
              _lib.ptr(self.out), B, h, w)
        self.op_info.append(("dc_conv7", 2.0 * B * h * w * 2 * 32 * 9))
        self.ops.append(lambda st: _lib.check(lib.pwc_conv3x3_flow_head_fwd(*ha, st), "dc_conv7"))
        self.launches = len(self.ops) + sum(lib.pwc_tc_conv_num_launches(L.handle) - 1 for L in self.layers.values() if L.handle)
