"""bf16-conv execution plan of the B200 PWC-Net forward (PWCDCNet.forward, /root/reference/models/PWCNet.py:180-273).

Same buffer plan as engine.Plan (one NHWC concat buffer per level, no torch.cat), but activations are stored in bf16
and every 3x3 convolution except conv1a runs on the 5th-gen tensor cores (csrc/conv_tc.cu: tcgen05.mma, TMEM
accumulators, TMA operands, fp32 accumulate, fused bias + LeakyReLU, output written into its concat slice).
conv1a (Cin = 3) and the 2-channel transposed convolutions are HBM-bound CUDA-core kernels (csrc/misc_bf16.cu); the
flow tensors (predict_flowL outputs, up_flow read by the warp, final flow2) stay fp32.  The whole forward is a fixed
launch list and is captured once into a CUDA graph.
"""
from __future__ import annotations

import ctypes
import os

import torch

from . import _lib
from .ops import pick_kc
from . import split as S
from .engine import (weights_key, C1_OFF, CORR_OFF, DC, DEC_START, DEC_WRITE, FLOW_SCALE, PYR_CH, UPFEAT_OFF, UPFLOW_OFF, chan_map,
                     concat_width)

BF = torch.bfloat16


def _roundup(a, b):
    return (a + b - 1) // b * b


S2D = {} if os.environ.get("PWC_S2D", "1") == "0" else {1: 4, 2: 2}            # pixels per super-pixel for the 16- and 32-channel pyramid levels (64 channels per row)
PYR_CP = [4] + [_roundup(c, 16) for c in PYR_CH[1:]]      # 16, 32, 64, 96, 128, 208


def _s2d_weights(name, g):
    """Thin layers (C = 16 or 32 channels) as a full-width GEMM: g adjacent pixels of a dense NHWC row are ONE
    super-pixel with g*C channels (a free view of the same memory), so every 128-byte operand row carries real data.
    The 3x3 conv becomes a 3x3 conv on [H][W/g] whose weight is block-structured:
      W'[(dx,co), (dx',ci), ky, KX] = W[co, ci, ky, kx],  kx = g*(KX-1) + dx' - dx + 1  when 0 <= kx <= 2, else 0
    (input pixel g*(X+KX-1)+dx' = g*X+dx + kx-1).  Same products, zeros added: results are unchanged."""
    def make(sd):
        w, b = name(sd) if callable(name) else (sd[name + ".weight"], sd[name + ".bias"])
        co, ci = w.shape[0], w.shape[1]
        w2 = torch.zeros(g, co, g, ci, 3, 3, device=w.device, dtype=torch.float32)
        for dx in range(g):
            for dxp in range(g):
                for KX in range(3):
                    kx = g * (KX - 1) + dxp - dx + 1
                    if 0 <= kx <= 2:
                        w2[dx, :, dxp, :, :, KX] = w[:, :, :, kx]
        return w2.reshape(g * co, g * ci, 3, 3), b.repeat(g)
    return make


def _stacked(wfn_or_key):
    """Tap-stacked weight tensor for layers with 9*Cout <= 128 (include/pwc_b200.h, out_f32 | 16): row tap*Cout + co of
    a [128, Cin, 3, 3] tensor holds W[co, :, tap//3, tap%3] at kernel position (0,0)."""
    def make(sd):
        w, b = wfn_or_key(sd) if callable(wfn_or_key) else (sd[wfn_or_key + ".weight"], sd[wfn_or_key + ".bias"])
        cout, cin = w.shape[0], w.shape[1]
        ws = torch.zeros(128, cin, 3, 3, device=w.device, dtype=torch.float32)
        for tap in range(9):
            ws[tap * cout:(tap + 1) * cout, :, 0, 0] = w[:, :, tap // 3, tap % 3]
        return ws, b
    return make


def _lanes(wfn_or_key, cout, head=None):
    """Tap-lane weight tensor for layers with Cout <= 64 (include/pwc_b200.h, out_f32 | 32; WmParams::tl): a [128, Cin, 3, 3]
    tensor whose packer tap slot t = 3*a + kx holds, for Cout <= 32 (three lanes of 32 rows), row 32*l + co = W[co, :, ky=l, kx] in
    slots a = 0; for Cout <= 64 (two lanes of 64 rows), rows co / 64 + co = W[co, :, ky = 0 / 1, kx] in slots a = 0 and
    row co = W[co, :, ky=2, kx] in slots a = 1.  Unused slots / rows stay zero and are never loaded.
    head = (wfn, nhead) (Cout <= 32 only): rows 96 + ky * nhead + h = kernel row ky of the nhead output channels of another layer
    over the same input channels (pwc_tc_conv_set_head_rows)."""
    def make(sd):
        w, b = wfn_or_key(sd) if callable(wfn_or_key) else (sd[wfn_or_key + ".weight"], sd[wfn_or_key + ".bias"])
        cin = w.shape[1]
        wl = torch.zeros(128, cin, 3, 3, device=w.device, dtype=torch.float32)
        if cout <= 32:
            for l in range(3):
                wl[32 * l:32 * l + cout, :, 0, :] = w[:, :, l, :]
            if head is not None:
                wh, nh = head[0](sd)[0], head[1]
                assert wh.shape[0] == nh and wh.shape[1] == cin and 3 * nh <= 32, (tuple(wh.shape), nh, cin)
                for ky in range(3):
                    wl[96 + ky * nh:96 + (ky + 1) * nh, :, 0, :] = wh[:, :, ky, :]
        else:
            wl[0:cout, :, 0, :] = w[:, :, 0, :]
            wl[64:64 + cout, :, 0, :] = w[:, :, 1, :]
            wl[0:cout, :, 1, :] = w[:, :, 2, :]
        return wl, b
    return make


TL_MODE = os.environ.get("PWC_TC_TL", "1") != "0"
TS_MODE = os.environ.get("PWC_TC_TS", "1") != "0"
S2D_DEC = os.environ.get("PWC_TC_S2D_DEC", "1") != "0"
# measured (B200, batch 8, 1024x448): the super-pixel form halves the MMA count but streams 4x larger weight tiles per item,
# so it only pays where items are plentiful (level 2) and the M operand was 3/4 empty (Cout = 32): conv2_4 178 -> 153 us
S2D_DEC_MAXLVL = int(os.environ.get("PWC_TC_S2D_DEC_MAXLVL", "2"))
S2D_DEC_MAXCOUT = int(os.environ.get("PWC_TC_S2D_DEC_MAXCOUT", "32"))
TL_MAXLVL = int(os.environ.get("PWC_TC_TL_MAXLVL", "3"))        # decoder levels whose Cout <= 64 layers run in tap-lane form
HEAD_ROWS = os.environ.get("PWC_TC_HEAD_ROWS", "1") != "0"       # predict_flow2's shared input channels as head rows of conv2_4


def _pfup_weights(lvl):
    """predict_flowL [2,Cin,3,3] stacked with upfeatL (ConvTranspose2d(Cin,2,4,2,1), PWCNet.py:35-36) rewritten as a 3x3
    convolution with 8 outputs: channel 2 + 2*(2a+b) + j is output j at pixel (2y+a, 2x+b);
    tap (dy,dx) of phase (a,b) uses transposed-conv tap (ky,kx) = (a+1-2dy, b+1-2dx) when inside [0,3]."""
    def make(sd):
        wp, bp = sd[f"predict_flow{lvl}.weight"], sd[f"predict_flow{lvl}.bias"]
        wt, bt = sd[f"upfeat{lvl}.weight"], sd[f"upfeat{lvl}.bias"]          # [Cin, 2, 4, 4]
        cin = wp.shape[1]
        w = torch.zeros(10, cin, 3, 3, device=wp.device, dtype=torch.float32)
        w[0:2] = wp
        for a in range(2):
            for b in range(2):
                for dy in (-1, 0, 1):
                    ky = a + 1 - 2 * dy
                    if not 0 <= ky <= 3:
                        continue
                    for dx in (-1, 0, 1):
                        kx = b + 1 - 2 * dx
                        if not 0 <= kx <= 3:
                            continue
                        c0 = 2 + 2 * (2 * a + b)
                        w[c0:c0 + 2, :, dy + 1, dx + 1] = wt[:, :, ky, kx].t()
        return w, torch.cat([bp, bt, bt, bt, bt])
    return make


# Layers routed to the warp-MMA small-level kernel (csrc/conv_mma_small.cu: pixel tile x channel slice x cluster K-split)
# instead of the tcgen05 kernels.  Measured on B200 (batch 8, 1024x448; ncu): conv6_4 18.9 -> 9.6 us, conv6_3 17.1 ->
# 12.6 us, conv6a / conv6b 31.3 -> 24.5 us; level-5 layers and the short-K level-6 layers are faster on tcgen05.
_SMALL_DEFAULT = "conv6_3.0,conv6_4.0,conv6a.0,conv6b.0"
SMALL_LAYERS = set(x for x in os.environ.get("PWC_MMA_SMALL_LAYERS", _SMALL_DEFAULT).split(",") if x and x != "none")


# Thin pyramid layers routed to the persistent warp-MMA kernel (csrc/conv_thin.cu).  Measured on B200 (batch 8): conv2a
# (16 -> 32, stride 2) 62.7 -> 41.6 us, conv2b 44 -> 41 us; the other stride-1 layers are on par with or slower than the tcgen05
# space-to-depth path (conv1aa 60.8 vs 58.7 us, conv3aa 37 vs 23 us).
THIN_LAYERS = set(x for x in os.environ.get("PWC_THIN_LAYERS", "conv1aa.0,conv1b.0,conv2a.0,conv2b.0").split(",") if x and x != "none")


class _TcLayer:
    __slots__ = ("key", "w", "b", "handle", "cmap", "cin", "cout", "span", "kc", "coutp", "wfn", "wrows", "groups", "kkskip",
                 "masks_set", "split_in", "kpad", "mask_key", "fused", "products")


class PlanBF16:
    def __init__(self, model, B, H, W, device, align_corners=False, use_graph=True, video=False, half="bf16"):
        """half: the 16-bit storage format of activations and weights, 'bf16' (conv_mode 'bf16') or 'f16' (conv_mode 'f16':
        IEEE half, 3 more mantissa bits; same kernels -- tcgen05 kind::f16 and mma.sync take either -- same speed).
        video=True: the B pairs are CONSECUTIVE frames (t, t+1) of one stream: the input is B+1 frames [B+1,3,H,W], the
        feature pyramid runs ONCE per frame (B+1 images instead of 2B) and pair k reads frame k+1's features in place --
        image k+1 of the same concat buffer -- as its image-2 features."""
        if H % 64 or W % 64:
            raise ValueError(f"H and W must be multiples of 64 (6 stride-2 levels), got {H}x{W}")
        self.lib = _lib.load()
        self.B, self.H, self.W, self.device = B, H, W, device
        self.align_corners = int(bool(align_corners))
        self.use_graph = use_graph
        if half not in ("bf16", "f16"):
            raise ValueError("half must be 'bf16' or 'f16'")
        self.half_fmt = 1 if half == "f16" else 0
        self.hdt = torch.float16 if half == "f16" else BF
        zb = lambda *s: torch.zeros(*s, device=device, dtype=self.hdt)      # noqa: E731
        zf = lambda *s: torch.zeros(*s, device=device, dtype=torch.float32)  # noqa: E731
        self.video = bool(video)
        NP = self.NP = (B + 1) if video else 2 * B            # images through the pyramid
        NC = (B + 1) if video else B                          # images holding pyramid features in the concat buffers
        self.x_static = zf(B + 1, 3, H, W) if video else zf(B, 6, H, W)
        self.hw = {l: (H >> l, W >> l) for l in range(1, 7)}
        self.pa = {l: zb(NP, *self.hw[l], PYR_CP[l]) for l in range(1, 7)}
        self.pb = {l: zb(NP, *self.hw[l], PYR_CP[l]) for l in range(1, 7)}
        self.cat, self.f2, self.flow, self.upflow = {}, {}, {}, {}
        for l in range(2, 7):
            h, w = self.hw[l]
            self.cat[l] = zb(NC, h, w, concat_width(l))
            self.f2[l] = None if video else zb(B, h, w, PYR_CP[l])
            self.flow[l] = zf(B, h, w, 4)
            if l < 6:
                self.upflow[l] = zf(B, h, w, 2)
        self.f1_6 = zb(NC, *self.hw[6], PYR_CP[6])
        h2, w2 = self.hw[2]
        self.dc = [zb(B, h2, w2, c) for (_, c, _) in DC]
        self.flow_out4 = zf(B, h2, w2, 4)
        self.out = zf(B, 2, h2, w2)
        self.weights_version = None
        self.layers = {}
        self.small = {}
        self.ops = []
        self.op_info = []          # (name, algorithmic flops) per launch, for profiling tools
        self.graph = None
        self.graph_tail = None     # ops[1:] as a graph, for run_from()
        self.launches = 0

    class _Fmt:
        """Selects this plan's 16-bit storage format for the library calls made inside the block (pwc_set_half_format is
        thread-local and read at create / launch time) and restores bf16, the library default, afterwards."""

        def __init__(self, plan):
            self.plan = plan

        def __enter__(self):
            if self.plan.half_fmt:
                _lib.check(self.plan.lib.pwc_set_half_format(self.plan.half_fmt), "pwc_set_half_format")

        def __exit__(self, *exc):
            if self.plan.half_fmt:
                self.plan.lib.pwc_set_half_format(0)
            return False

    def __del__(self):
        try:
            for L in self.layers.values():
                if L.handle:
                    self.lib.pwc_tc_conv_destroy(L.handle)
        except Exception:  # noqa: BLE001
            pass

    # ------------------------------------------------------------------ layer construction
    def _tc(self, key, cin, cout, src, src_coff, span, dst, dst_coff, N, hin, win, stride=1, dil=1, slope=0.1,
            cmap=None, src2=None, src2_coff=0, dst2=None, dst2_coff=0, n_split=0, out_f32=False, residual=None, wfn=None,
            flops=None, ts=False, groups=None, kkskip=False, in_lo=0, out_lo=0, up_lo=0, up_dup=0, kpad=None, no_small=False,
            tl=False, head=None, partial=None):
        """in_lo / out_lo: channel offset of the lo plane of a split-precision input / output (0 = single precision);
        up_lo / up_dup: lo / second-hi offsets of the fused upfeat outputs; kpad = (idx, lo_pos, dup_pos): the torch input
        channels idx are also read from their lo copies / second hi copies at those slice positions (split.kpad_weights)."""
        if key in THIN_LAYERS and dil == 1 and not out_f32 and not ts and groups is None and residual is None and cmap is None \
                and wfn is None and cin == span and cin in (16, 32, 64) and cout in (16, 32, 64) and not (stride == 2 and cin > 16) \
                and not (in_lo or out_lo):
            return self._thin(key, cin, cout, src, src_coff, dst, dst_coff, N, hin, win, stride, slope, src2, src2_coff, dst2,
                              dst2_coff, n_split)
        if key in SMALL_LAYERS and dil == 1 and not out_f32 and not ts and groups is None and residual is None \
                and (dst2 is None or dst2_coff == 0) and not (in_lo or out_lo or no_small):
            return self._small(key, cin, cout, src, src_coff, span, dst, dst_coff, N, hin, win, stride, slope, cmap, src2,
                               src2_coff, dst2, dst2_coff, n_split, wfn, flops)
        L = _TcLayer()
        L.kkskip = bool(kkskip and span <= 64 and cmap is None and os.environ.get("PWC_TC_KKSKIP", "1") != "0")
        ts = bool(ts and TS_MODE and 9 * cout <= 128 and stride == 1 and dil == 1 and out_f32)
        tl = bool(tl and TL_MODE and cout <= 64 and stride == 1 and dil == 1 and not out_f32 and not ts and groups is None
                  and dst2 is None and src2 is None)
        # head = (weight fn, nhead, fp32 buffer [N,H,W,nhead]): extra output channels in the idle lane quarter of a three-lane layer;
        # partial = such a buffer, added by a tap-stacked layer before its activation (include/pwc_b200.h)
        if head is not None and not (tl and cout <= 32):
            raise RuntimeError(f"{key}: head rows need the three-lane tap-lane form")
        if partial is not None and not ts:
            raise RuntimeError(f"{key}: a partial input needs the tap-stacked form")
        L.wfn = _stacked(wfn if wfn is not None else key) if ts else \
            (_lanes(wfn if wfn is not None else key, cout, head=None if head is None else head[:2]) if tl else wfn)
        L.key, L.cin, L.cout, L.span, L.cmap = key, cin, cout, span, cmap
        L.coutp = 128 if (ts or tl) else _roundup(cout, 16)
        L.kc = pick_kc(span, stride, dil, src2 is not None, W=win, coutp=L.coutp, out_f32=bool(out_f32) or dst2 is not None)
        L.wrows = 128 if (ts or tl) else cout           # rows of the weight tensor handed to the packer
        L.split_in, L.kpad, L.mask_key = bool(in_lo), kpad, key
        L.products = 2 if in_lo < 0 else 3                # in_lo < 0: split activations only (lo plane at -in_lo)
        # adjacent [hi | lo] planes that fit one swizzle row run the fused form of the stride-2 kernel (TcConvParams::fused)
        L.fused = in_lo > 0 and stride == 2 and in_lo == span == L.kc and L.kc <= 32 and cmap is None and \
            os.environ.get("PWC_TC_FUSED_SPLIT", "1") != "0"
        if out_lo and not tl:
            L.coutp = _roundup(cout, 32)                  # split stores exist in the per-warp (32-channel group) store path
        dev = self.device
        L.w = torch.zeros(9 * _roundup(span, L.kc) * L.coutp * ((4 if L.fused else L.products) if in_lo else 1), device=dev, dtype=self.hdt)
        L.b = torch.zeros(cout, device=dev, dtype=torch.float32)
        h = ctypes.c_void_p()
        args = (ctypes.byref(h), _lib.ptr(src), src.shape[-1], src_coff,
                _lib.ptr(src2), (src2.shape[-1] if src2 is not None else 0), src2_coff, n_split, span, L.kc,
                _lib.ptr(L.w), _lib.ptr(L.b), cout, L.coutp, _lib.ptr(dst), dst.shape[-1], dst_coff,
                _lib.ptr(dst2), (dst2.shape[-1] if dst2 is not None else 0), dst2_coff, int(out_f32) | (16 if ts else 0) | (32 if tl else 0),
                _lib.ptr(residual), (residual.shape[-1] if residual is not None else 0), N, hin, win, stride, dil, slope)
        if in_lo or out_lo or up_lo or up_dup:
            e = self.lib.pwc_tc_conv_create_split(*args, in_lo, out_lo, up_lo, up_dup)
        else:
            e = self.lib.pwc_tc_conv_create(*args)
        _lib.check(e, f"pwc_tc_conv_create {key}")
        if head is not None:
            _lib.check(self.lib.pwc_tc_conv_set_head_rows(h, _lib.ptr(head[2]), head[1]), f"set_head_rows {key}")
        if partial is not None:
            _lib.check(self.lib.pwc_tc_conv_set_partial(h, _lib.ptr(partial), partial.shape[-1]), f"set_partial {key}")
        L.groups = groups
        if groups is not None and groups[1] != L.coutp // groups[0]:
            _lib.check(self.lib.pwc_tc_conv_set_out_groups(h, groups[0], groups[1]), f"set_out_groups {key}")
        L.handle = h
        self.layers[key] = L
        lib = self.lib
        ho, wo = (hin - 1) // stride + 1, (win - 1) // stride + 1
        self.op_info.append((key, flops if flops is not None else 2.0 * N * ho * wo * cout * cin * 9))
        self.ops.append(lambda st: _lib.check(lib.pwc_tc_conv_launch(h, st), key))

    def _thin(self, key, cin, cout, src, src_coff, dst, dst_coff, N, hin, win, stride, slope, src2, src2_coff, dst2, dst2_coff,
              n_split):
        """Same layer contract as _tc, launched through pwc_conv3x3_thin_fwd (weights [9][cout][cin])."""
        L = _TcLayer()
        L.kkskip, L.groups, L.handle, L.wfn, L.cmap = False, None, None, None, None
        L.key, L.cin, L.cout, L.span, L.coutp, L.kc, L.wrows = key, cin, cout, cin, cout, cin, cout
        dev = self.device
        L.w = torch.zeros(9 * cin * cout, device=dev, dtype=self.hdt)
        L.b = torch.zeros(cout, device=dev, dtype=torch.float32)
        self.layers[key] = L
        lib = self.lib
        n_in = n_split if src2 is not None else N
        n_out = n_split if dst2 is not None else N
        args = (_lib.ptr(src), src.shape[-1], src_coff, _lib.ptr(src2), (src2.shape[-1] if src2 is not None else 0), src2_coff,
                n_in, cin, _lib.ptr(L.w), _lib.ptr(L.b), cout, _lib.ptr(dst), dst.shape[-1], dst_coff,
                _lib.ptr(dst2), (dst2.shape[-1] if dst2 is not None else 0), dst2_coff, n_out, N, hin, win, stride, slope)
        ho, wo = (hin - 1) // stride + 1, (win - 1) // stride + 1
        self.op_info.append((key, 2.0 * N * ho * wo * cout * cin * 9))
        self.ops.append(lambda st: _lib.check(lib.pwc_conv3x3_thin_fwd(*args, st), key))

    def _small(self, key, cin, cout, src, src_coff, span, dst, dst_coff, N, hin, win, stride, slope, cmap, src2, src2_coff,
               dst2, dst2_coff, n_split, wfn, flops):
        """Same layer contract as _tc, launched through pwc_conv3x3_mma_fwd (weights packed with kc = 64)."""
        L = _TcLayer()
        L.kkskip, L.groups, L.handle = False, None, None
        L.wfn, L.key, L.cin, L.cout, L.span, L.cmap = wfn, key, cin, cout, span, cmap
        L.coutp, L.kc, L.wrows = _roundup(cout, 16), 64, cout
        dev = self.device
        L.w = torch.zeros(9 * _roundup(span, 64) * L.coutp, device=dev, dtype=self.hdt)
        L.b = torch.zeros(cout, device=dev, dtype=torch.float32)
        self.layers[key] = L
        lib = self.lib
        n_in = n_split if src2 is not None else N
        n_out = n_split if dst2 is not None else N
        args = (_lib.ptr(src), src.shape[-1], src_coff, _lib.ptr(src2), (src2.shape[-1] if src2 is not None else 0), src2_coff,
                n_in, span, _lib.ptr(L.w), _lib.ptr(L.b), cout, L.coutp, _lib.ptr(dst), dst.shape[-1], dst_coff,
                _lib.ptr(dst2), (dst2.shape[-1] if dst2 is not None else 0), dst2_coff, n_out, N, hin, win, stride, slope)
        ho, wo = (hin - 1) // stride + 1, (win - 1) // stride + 1
        self.op_info.append((key, flops if flops is not None else 2.0 * N * ho * wo * cout * cin * 9))
        self.ops.append(lambda st: _lib.check(lib.pwc_conv3x3_mma_fwd(*args, st), key))

    def _deconv(self, key, cin, src, src_coff, span, dst, dst_coff, hin, win, dst_b=None):
        dev = self.device
        cinp = _roundup(span, 4)
        w = torch.zeros(32 * cinp, device=dev, dtype=torch.float32)
        b = torch.zeros(2, device=dev, dtype=torch.float32)
        self.small[key] = (w, b, cin, cinp)
        lib, B = self.lib, self.B
        idt = _lib.BF16 if src.dtype in (BF, torch.float16) else _lib.F32      # BF16 = "the 16-bit storage format"
        odt = _lib.BF16 if dst.dtype in (BF, torch.float16) else _lib.F32
        args = (_lib.ptr(src), idt, src.shape[-1], src_coff, span, _lib.ptr(w), _lib.ptr(b), _lib.ptr(dst), odt,
                dst.shape[-1], dst_coff, _lib.ptr(dst_b), (dst_b.shape[-1] if dst_b is not None else 0), B, hin, win)
        self.op_info.append((key, 2.0 * B * hin * win * 4 * 2 * cin * 4))
        self.ops.append(lambda st: _lib.check(lib.pwc_deconv4x4s2_c2_fwd(*args, st), key))

    def _corr(self, lvl):
        h, w = self.hw[lvl]
        cat = self.cat[lvl]
        C = PYR_CH[lvl]
        if lvl == 6:
            f1p, f1_pitch, flow, fp, scale = _lib.ptr(self.f1_6), PYR_CP[6], None, 0, 1.0
        else:
            f1p = ctypes.c_void_p(cat.data_ptr() + 2 * C1_OFF)
            f1_pitch, flow, fp, scale = cat.shape[-1], self.upflow[lvl], 2, FLOW_SCALE[lvl]
        outp = ctypes.c_void_p(cat.data_ptr() + 2 * CORR_OFF)
        lib = self.lib
        f2p, f2_pitch = _lib.ptr(self.f2[lvl]), PYR_CP[lvl]
        if self.video:                      # image-2 features of pair k = image-1 features of frame k+1, read in place
            if lvl == 6:
                f2p = ctypes.c_void_p(self.f1_6.data_ptr() + 2 * h * w * PYR_CP[6])
            else:
                f2p, f2_pitch = ctypes.c_void_p(cat.data_ptr() + 2 * h * w * cat.shape[-1] + 2 * C1_OFF), cat.shape[-1]
        args = (f1p, f1_pitch, f2p, f2_pitch, _lib.ptr(flow), fp, scale, self.align_corners,
                0.9999, 0.1, outp, cat.shape[-1], self.B, C, h, w, 4, _lib.NHWC, _lib.BF16)
        self.op_info.append((f"warp_corr{lvl}", 2.0 * self.B * h * w * 81 * C))
        self.ops.append(lambda st: _lib.check(lib.pwc_warp_corr_lrelu_fwd(*args, st), f"warp_corr{lvl}"))

    def _build(self):
        B, lib = self.B, self.lib
        dev = self.device
        # conv1a on CUDA cores, straight from the NCHW fp32 input
        self.w1a = torch.zeros(16, 3, 3, 3, device=dev)
        self.b1a = torch.zeros(16, device=dev)
        NP, vid = self.NP, self.video
        a1 = (_lib.ptr(self.x_static), _lib.ptr(self.w1a), _lib.ptr(self.b1a), _lib.ptr(self.pa[1]), (NP if vid else B), self.H,
              self.W, 0.1)
        self.op_info.append(("conv1a.0", 2.0 * NP * (self.H // 2) * (self.W // 2) * 16 * 27))
        c1a = lib.pwc_conv1a_frames_bf16_fwd if vid else lib.pwc_conv1a_bf16_fwd
        self.ops.append(lambda st: _lib.check(c1a(*a1, st), "conv1a"))
        self._conv1a_on = lambda x, st: _lib.check(c1a(_lib.ptr(x), *a1[1:], st), "conv1a")      # ops[0] on another input
        for l in range(1, 7):
            names = (f"conv{l}a", f"conv{l}aa", f"conv{l}b") if l < 6 else ("conv6aa", "conv6a", "conv6b")
            h, w = self.hw[l]
            if l > 1:
                hin, win = self.hw[l - 1]
                cin = PYR_CH[l - 1]
                if l == 2:
                    self._tc(names[0] + ".0", cin, PYR_CH[l], self.pa[1], 0, PYR_CP[1], self.pa[l], 0, NP, hin, win, stride=2)
                elif vid:
                    self._tc(names[0] + ".0", cin, PYR_CH[l], self.cat[l - 1], C1_OFF, PYR_CP[l - 1], self.pa[l], 0, NP,
                             hin, win, stride=2)
                else:
                    self._tc(names[0] + ".0", cin, PYR_CH[l], self.cat[l - 1], C1_OFF, PYR_CP[l - 1], self.pa[l], 0, NP,
                             hin, win, stride=2, src2=self.f2[l - 1], n_split=B)
            g = S2D.get(l, 1)            # thin dense->dense layers run on the space-to-depth view (see _s2d_weights)
            C = PYR_CH[l]
            fl = 2.0 * NP * h * w * C * C * 9
            va = self.pa[l].view(NP, h, w // g, g * PYR_CP[l]) if g > 1 else self.pa[l]
            vb = self.pb[l].view(NP, h, w // g, g * PYR_CP[l]) if g > 1 else self.pb[l]
            if g > 1 and names[1] + ".0" not in THIN_LAYERS:
                self._tc(names[1] + ".0", g * C, g * C, va, 0, g * C, vb, 0, NP, h, w // g,
                         wfn=_s2d_weights(names[1] + ".0", g), flops=fl, kkskip=True)
            else:
                self._tc(names[1] + ".0", C, C, self.pa[l], 0, PYR_CP[l], self.pb[l], 0, NP, h, w)
            if l == 1 and names[2] + ".0" in THIN_LAYERS:
                self._tc(names[2] + ".0", C, C, self.pb[l], 0, PYR_CP[l], self.pa[l], 0, NP, h, w)
            elif l == 1:
                self._tc(names[2] + ".0", g * C, g * C, vb, 0, g * C, va, 0, NP, h, w // g,
                         wfn=_s2d_weights(names[2] + ".0", g), flops=fl, kkskip=True)
            elif l < 6:
                self._tc(names[2] + ".0", PYR_CH[l], PYR_CH[l], self.pb[l], 0, PYR_CP[l], self.cat[l], C1_OFF, NP, h, w,
                         dst2=self.f2[l], n_split=(NP if vid else B))
            else:
                self._tc(names[2] + ".0", PYR_CH[l], PYR_CH[l], self.pb[l], 0, PYR_CP[l], self.f1_6, 0, NP, h, w,
                         dst2=self.f2[l], n_split=(NP if vid else B))
        self.head_part = {}
        for l in (6, 5, 4, 3, 2):
            h, w = self.hw[l]
            cat, wd = self.cat[l], concat_width(l)
            nin = 81 + (0 if l == 6 else PYR_CH[l] + 4)
            self._corr(l)
            for k in range(5):
                cm = chan_map(l, DEC_START[k])
                cout = [128, 128, 96, 64, 32][k]
                g = 128 // cout
                if TL_MODE and cout <= 64 and l <= TL_MAXLVL:
                    # Cout = 64 / 32: tap lanes -- 6 / 3 MMA groups per chunk instead of 9 (WmParams::tl)
                    head = None
                    if k == 4 and l == 2 and HEAD_ROWS and DEC_START[4] == 32:
                        # predict_flow2's sums over the 533 channels this layer reads ride in its idle MMA rows (include/pwc_b200.h,
                        # pwc_tc_conv_set_head_rows): the head's own pass over the concat buffer cost 69 us for 2 output channels
                        hsuf = (lambda sd: (sd["predict_flow2.weight"][:, 32:].contiguous(), sd["predict_flow2.bias"]))
                        self.head_part[l] = torch.zeros(B, h, w, 2, device=dev, dtype=torch.float32)
                        head = (hsuf, 2, self.head_part[l])
                    self._tc(f"conv{l}_{k}.0", len(cm), cout, cat, DEC_START[k], wd - DEC_START[k], cat, DEC_WRITE[k], B, h, w,
                             cmap=cm, tl=True, head=head)
                    continue
                if S2D_DEC and cout <= S2D_DEC_MAXCOUT and l <= S2D_DEC_MAXLVL and w % g == 0:
                    # Cout = 64 / 32: g adjacent pixels form one super-pixel so that all 128 MMA rows carry outputs
                    # (weight block-structured as in _s2d_weights; all-zero (tap, chunk) blocks are skipped)
                    v = cat[:B].view(B, h, w // g, g * wd)
                    cm2 = [dxp * wd + DEC_START[k] + c for dxp in range(g) for c in cm]
                    self._tc(f"conv{l}_{k}.0", g * len(cm), g * cout, v, 0, g * wd, v, DEC_WRITE[k], B, h, w // g, cmap=cm2,
                             wfn=_s2d_weights(f"conv{l}_{k}.0", g), flops=2.0 * B * h * w * cout * len(cm) * 9,
                             groups=(g, wd))
                    continue
                self._tc(f"conv{l}_{k}.0", len(cm), cout, cat, DEC_START[k], wd - DEC_START[k], cat,
                         DEC_WRITE[k], B, h, w, cmap=cm)
            cm = chan_map(l, 0)
            if l > 2:
                # predict_flowL and upfeatL read the same tensor: ONE conv with 2 + 8 output channels (the 4x4 stride-2
                # transposed conv is 4 phase-shifted 3x3 convs); the epilogue scatters the 8 to the 2x-upsampled slice
                self._tc(f"pfup{l}", len(cm), 10, cat, 0, wd, self.flow[l], 0, B, h, w, slope=1.0, cmap=cm, out_f32=2,
                         dst2=self.cat[l - 1], dst2_coff=UPFEAT_OFF, wfn=_pfup_weights(l), ts=True)
                self._deconv(f"deconv{l}", 2, self.flow[l], 0, 4, self.cat[l - 1], UPFLOW_OFF, h, w, dst_b=self.upflow[l - 1])
            elif l in self.head_part:
                first32 = (lambda sd: (sd["predict_flow2.weight"][:, :32].contiguous(), sd["predict_flow2.bias"]))
                self._tc(f"predict_flow{l}", 32, 2, cat, 0, 32, self.flow[l], 0, B, h, w, slope=1.0, out_f32=True, ts=True,
                         wfn=first32, flops=2.0 * B * h * w * 2 * len(cm) * 9, partial=self.head_part[l])
            else:
                self._tc(f"predict_flow{l}", len(cm), 2, cat, 0, wd, self.flow[l], 0, B, h, w, slope=1.0, cmap=cm, out_f32=True,
                         ts=True)
        h, w = self.hw[2]
        cm = chan_map(2, 0)
        src, cin, span, cmap = self.cat[2], len(cm), concat_width(2), cm
        for i, (name, cout, dil) in enumerate(DC):
            g = 128 // cout if cout <= S2D_DEC_MAXCOUT else 1
            if TL_MODE and cout <= 64 and dil == 1 and span >= 128:      # (dc_conv6, one chunk: 28 us as s2d, 49 us in lanes)
                self._tc(name + ".0", cin, cout, src, 0, span, self.dc[i], 0, B, h, w, cmap=cmap, tl=True)
            elif S2D_DEC and g > 1 and dil == 1 and cmap is None and w % g == 0 and src.shape[-1] == span:
                vi, vo = src[:B].view(B, h, w // g, g * span), self.dc[i].view(B, h, w // g, g * cout)
                self._tc(name + ".0", g * cin, g * cout, vi, 0, g * span, vo, 0, B, h, w // g,
                         wfn=_s2d_weights(name + ".0", g), flops=2.0 * B * h * w * cout * cin * 9, groups=(g, cout))
            else:
                self._tc(name + ".0", cin, cout, src, 0, span, self.dc[i], 0, B, h, w, dil=dil, cmap=cmap)
            src, cin, span, cmap = self.dc[i], cout, cout, None
        if os.environ.get("PWC_FLOW_HEAD", "1") != "0":
            # dc_conv7 + residual + NCHW unpack in ONE launch of the thin-layer kernel (csrc/conv_thin.cu, flow-head mode):
            # 32 + 4 us of tap-stacked tcgen05 conv and unpack kernel -> 12 us
            L = _TcLayer()
            L.kkskip, L.groups, L.handle, L.cmap = False, None, None, None
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
        else:
            self._tc("dc_conv7", 32, 2, src, 0, 32, self.flow_out4, 0, B, h, w, slope=1.0, out_f32=True, residual=self.flow[2],
                     ts=True)
            uo = (_lib.ptr(self.flow_out4), 4, 0, _lib.ptr(self.out), B, h, w, 1.0)
            self.op_info.append(("unpack_flow", 0.0))
            self.ops.append(lambda st: _lib.check(lib.pwc_unpack_flow_f32(*uo, st), "unpack_flow"))
        self.launches = len(self.ops) + sum(self.lib.pwc_tc_conv_num_launches(L.handle) - 1 for L in self.layers.values() if L.handle)

    # ------------------------------------------------------------------ weights
    def pack_weights(self, model):
        ver = weights_key(model)
        if ver == self.weights_version:
            return
        if not self.ops:
            with self._Fmt(self):
                self._build()
        with self._Fmt(self):
            self._pack_weights(model, ver)

    def _pack_weights(self, model, ver):
        lib, dev = self.lib, self.device
        st = _lib.stream_handle()
        sd = {k: v.detach().to(device=dev, dtype=torch.float32).contiguous() for k, v in model.state_dict().items()}
        keep = []
        for key, L in self.layers.items():
            wt, bs = (sd[key + ".weight"], sd[key + ".bias"]) if L.wfn is None else L.wfn(sd)
            wt = wt.contiguous()
            assert wt.shape[0] == L.wrows and wt.shape[1] == L.cin, (key, tuple(wt.shape), L.cin, L.wrows)
            cm, span_pack, kc_pack = L.cmap, L.span, L.kc
            if getattr(L, "fused", False):
                # fused split: per tap two weight tiles of 2*cin columns, [W_hi | W_hi] and [W_lo | 0]
                wt = torch.cat([wt, wt, S.residual(wt, self.hdt)], 1).contiguous()
                span_pack, kc_pack = 4 * L.span, 2 * L.span
            elif getattr(L, "split_in", False):
                # split-precision input: three weight chunks per logical chunk (W_hi | W_lo | W_hi), see split.virtual_weights
                wt, cm, span_pack = S.virtual_weights(wt, self.hdt, L.kc, L.span, cm, L.products)
            mp = None
            if cm is not None:
                mp = torch.tensor(cm, device=dev, dtype=torch.int32); keep.append(mp)
            _lib.check(lib.pwc_pack_conv_weight_bf16(_lib.ptr(wt), _lib.ptr(L.w), _lib.ptr(mp), L.wrows, wt.shape[1], span_pack,
                                                     kc_pack, L.coutp, st), f"pack {key}")
            L.b.copy_(bs)
            if (L.kkskip or L.groups is not None) and not getattr(L, "masks_set", False):
                # Skip masks come from the STRUCTURE of the space-to-depth weight (which (tap, K block) pairs can ever be
                # non-zero: the weight builder applied to all-ones parameters), never from the values: they are kernel
                # parameters baked into the captured CUDA graph and must not change when a new state_dict is loaded.
                # (1 + 2^-12: its 16-bit rounding residual is non-zero too, so W_lo copies of a K-padded layer count as present)
                ones = {k: torch.full_like(v, 1.000244140625) for k, v in sd.items()}
                ws = (L.wfn(ones)[0] != 0)
                if L.kkskip:
                    # single-chunk layer: all-zero 16-channel K steps of each tap's weight tile are not issued
                    km = (ctypes.c_int * 9)()
                    nkk = (L.span + 15) // 16
                    for tap in range(9):
                        blk = ws[:, :, tap // 3, tap % 3]
                        for kk in range(nkk):
                            if not bool(blk[:, 16 * kk:16 * kk + 16].any()):
                                km[tap] |= 1 << kk
                    if all(km[t] == (1 << nkk) - 1 for t in range(9)):
                        km = (ctypes.c_int * 9)()                # nothing structurally non-zero: issue everything
                    _lib.check(lib.pwc_tc_conv_set_kk_skip(L.handle, km), f"set_kk_skip {key}")
                if L.groups is not None:
                    # zero-block skip list of the block-structured weight: (tap, 64-channel chunk) pairs
                    nchunk = (L.span + L.kc - 1) // L.kc
                    cm = torch.tensor(L.cmap, device=dev, dtype=torch.long) if L.cmap is not None else torch.arange(L.cin, device=dev)
                    masks = (ctypes.c_ulonglong * 9)()
                    for tap in range(9):
                        nz = ws[:, :, tap // 3, tap % 3].any(0)
                        used = torch.zeros(nchunk, dtype=torch.bool, device=dev)
                        used[(cm[nz] // L.kc).unique()] = True
                        m = 0
                        for ch in (~used).nonzero().flatten().tolist():
                            m |= 1 << ch
                        masks[tap] = m
                    _lib.check(lib.pwc_tc_conv_set_skip(L.handle, masks), f"set_skip {key}")
                L.masks_set = True
        for key, (w, b, cin, cinp) in self.small.items():
            wt = sd[key + ".weight"]
            mp = None
            if key.startswith("upfeat"):
                mp = torch.tensor(chan_map(int(key[-1]), 0), device=dev, dtype=torch.int32); keep.append(mp)
            _lib.check(lib.pwc_pack_deconv_weight_f32(_lib.ptr(wt), _lib.ptr(w), _lib.ptr(mp), cin, cinp, st), f"pack {key}")
            b.copy_(sd[key + ".bias"])
        self.w1a.copy_(sd["conv1a.0.weight"]); self.b1a.copy_(sd["conv1a.0.bias"])
        torch.cuda.current_stream().synchronize()
        self.weights_version = ver

    # ------------------------------------------------------------------ forward
    def _launch_all(self):
        st = _lib.stream_handle()
        with self._Fmt(self):
            for op in self.ops:
                op(st)

    def run(self, x=None, all_levels=False):
        """x: [B,6,H,W] fp32 (device, or pinned host: it is copied into the plan's static input).  Returns flow2
        [B,2,H/4,W/4] fp32 (a plan-owned tensor)."""
        if x is not None and x.data_ptr() != self.x_static.data_ptr():
            self.x_static.copy_(x, non_blocking=True)
        if not self.use_graph:
            self._launch_all()
            return self.out
        if self.graph is None:
            self._launch_all()                      # warm-up outside capture (cudaFuncSetAttribute etc.)
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch_all()
            self.graph = g
        self.graph.replay()
        return self.out

    def run_from(self, x, release=None):
        """Same as run(), but the first layer reads the caller's device tensor `x` ([B,6,H,W] fp32, contiguous; [B+1,3,H,W]
        for a video plan) in place: conv1a -- the only reader of the input -- is launched eagerly on `x` and the other
        launches replay as one CUDA graph.  Saves the 2 x 11 MB/pair device-to-device copy into the static input; `x` may
        be overwritten as soon as conv1a has run: `release` (a torch.cuda.Event) is recorded right behind it."""
        if x.dtype != torch.float32 or not x.is_contiguous() or x.shape != self.x_static.shape or x.data_ptr() % 16:
            raise RuntimeError("run_from: a contiguous, 16-byte aligned fp32 tensor of the plan's input shape expected")
        if not self.use_graph:
            self.x_static.copy_(x, non_blocking=True)
            if release is not None:
                release.record(torch.cuda.current_stream())
            self._launch_all()
            return self.out
        if self.graph_tail is None:
            self._launch_all()
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g), self._Fmt(self):
                st = _lib.stream_handle()
                for op in self.ops[1:]:
                    op(st)
            self.graph_tail = g
        with self._Fmt(self):
            self._conv1a_on(x, _lib.stream_handle())
        if release is not None:
            release.record(torch.cuda.current_stream())
        self.graph_tail.replay()
        return self.out

    def time_ops(self, reps=3, inner=8):
        """GPU duration of every launch of the forward (profiling aid for bench.py / tools): each op is captured `inner`
        times back to back into its own CUDA graph and the replay is timed with CUDA events, so the host-side launch cost
        (~10 us per ctypes call, more than the small kernels themselves) stays out of the figure.  Every op is idempotent on
        the plan's static buffers.  Returns [(name, seconds, algorithmic flops)]."""
        st_cur = torch.cuda.current_stream()
        self._launch_all()
        st_cur.synchronize()
        out = []
        side = torch.cuda.Stream(self.device)
        for (name, fl), op in zip(self.op_info, self.ops):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side), self._Fmt(self):
                op(_lib.stream_handle())
                side.synchronize()
                with torch.cuda.graph(g, stream=side):
                    for _ in range(inner):
                        op(_lib.stream_handle())
                g.replay()
                side.synchronize()
                best = None
                for _ in range(reps):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(side); g.replay(); e1.record(side); e1.synchronize()
                    t = e0.elapsed_time(e1) * 1e-3 / inner
                    best = t if best is None else min(best, t)
            out.append((name, best, fl))
            del g
        return out

    def level_flow(self, lvl):
        return self.flow[lvl][..., :2].permute(0, 3, 1, 2).contiguous()
