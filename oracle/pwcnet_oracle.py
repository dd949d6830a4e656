"""CPU oracle for the PWC-Net (PWCDCNet) inference forward.  TEST INFRASTRUCTURE ONLY.

This file is a plain CPU restatement of the reference hot path.  Nothing under
``opticalflow_b200/`` may import it: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs use it, and only
as the checker / the timed CPU baseline.

What is restated, and where it lives in the reference (paths relative to
``/root/reference``):

* network topology, parameter names and init ........ models/PWCNet.py:45-138
* forward (pyramid, decoders, context net) ........... models/PWCNet.py:180-273
* warp (meshgrid + grid_sample + 0.9999 mask) ........ models/PWCNet.py:141-177
* cost volume (shift-and-dot) ........................ models/correlation_package/correlation.py:12-40
  with the CUDA kernel's normalisation ``acc / (k*k*C)``
  ..................................................... models/correlation_package/correlation_cuda_kernel.cu:104,143
  and its output size / channel order ................ correlation_cuda.cc:25-38, correlation_cuda_kernel.cu:107-110,139-143

The convolutions, transposed convolutions and ``grid_sample`` are third-party
arithmetic (PyTorch ATen, pip wheel ``torch`` -- the reference pins no version;
this image has torch 2.11.0).  They are executed through torch's CPU kernels,
exactly as the reference calls them.  A second, torch-free restatement of the
warp + cost volume arithmetic (``warp_corr_lrelu_numpy``) follows ATen's
``GridSampler.h`` formulas literally; tests check the two against each other.

Parity pinning: the reference has no tests or golden vectors,
so the oracle is pinned against the *unmodified reference module itself*,
imported in the build container by ``oracle/ref_harness.py``; the resulting
vectors are committed under ``tests/golden/`` by ``oracle/make_golden.py``.
"""
from __future__ import annotations

from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

MD = 4
ND = (2 * MD + 1) ** 2  # 81
DD = [128, 256, 352, 416, 448]  # np.cumsum([128,128,96,64,32]) -- PWCNet.py:75
PYR_CH = [3, 16, 32, 64, 96, 128, 196]  # PWCNet.py:52-69
FLOW_SCALE = {5: 0.625, 4: 1.25, 3: 2.5, 2: 5.0}  # PWCNet.py:212,226,240,256


def layer_specs():
    """Ordered list of (name, kind, cin, cout, k, stride, pad, dil) exactly in the
    order ``PWCDCNet.__init__`` registers its modules (PWCNet.py:52-132)."""
    L = []

    def conv(name, cin, cout, stride=1, pad=1, dil=1):
        L.append((name, "conv", cin, cout, 3, stride, pad, dil))

    def pflow(name, cin):
        L.append((name, "pred", cin, 2, 3, 1, 1, 1))

    def deconv(name, cin, cout):
        L.append((name, "deconv", cin, cout, 4, 2, 1, 1))

    conv("conv1a", 3, 16, 2); conv("conv1aa", 16, 16); conv("conv1b", 16, 16)
    conv("conv2a", 16, 32, 2); conv("conv2aa", 32, 32); conv("conv2b", 32, 32)
    conv("conv3a", 32, 64, 2); conv("conv3aa", 64, 64); conv("conv3b", 64, 64)
    conv("conv4a", 64, 96, 2); conv("conv4aa", 96, 96); conv("conv4b", 96, 96)
    conv("conv5a", 96, 128, 2); conv("conv5aa", 128, 128); conv("conv5b", 128, 128)
    conv("conv6aa", 128, 196, 2); conv("conv6a", 196, 196); conv("conv6b", 196, 196)
    for lvl, extra in ((6, 0), (5, 128 + 4), (4, 96 + 4), (3, 64 + 4), (2, 32 + 4)):
        od = ND + extra
        conv(f"conv{lvl}_0", od, 128)
        conv(f"conv{lvl}_1", od + DD[0], 128)
        conv(f"conv{lvl}_2", od + DD[1], 96)
        conv(f"conv{lvl}_3", od + DD[2], 64)
        conv(f"conv{lvl}_4", od + DD[3], 32)
        pflow(f"predict_flow{lvl}", od + DD[4])
        deconv(f"deconv{lvl}", 2, 2)
        if lvl > 2:
            deconv(f"upfeat{lvl}", od + DD[4], 2)
    od = ND + 32 + 4
    conv("dc_conv1", od + DD[4], 128, 1, 1, 1)
    conv("dc_conv2", 128, 128, 1, 2, 2)
    conv("dc_conv3", 128, 128, 1, 4, 4)
    conv("dc_conv4", 128, 96, 1, 8, 8)
    conv("dc_conv5", 96, 64, 1, 16, 16)
    conv("dc_conv6", 64, 32, 1, 1, 1)
    pflow("dc_conv7", 32)
    return L


def param_key(name, kind):
    """state_dict key prefix: ``conv()`` wraps the Conv2d in nn.Sequential -> '.0.'
    (PWCNet.py:26-30); predict_flow / deconv are bare modules (PWCNet.py:32-36)."""
    return f"{name}.0" if kind == "conv" else name


def reference_like_state_dict(seed=0):
    """Random-init weights bit-identical to ``torch.manual_seed(seed); PWCDCNet()``.

    Replays the reference constructor's RNG consumption: every module is built in
    registration order (each nn.Conv2d / nn.ConvTranspose2d draws its default init),
    then ``kaiming_normal_(mode='fan_in')`` + zero bias is applied in
    ``self.modules()`` order (PWCNet.py:134-138).
    """
    torch.manual_seed(seed)
    mods = OrderedDict()
    for (name, kind, cin, cout, k, s, p, d) in layer_specs():
        if kind == "deconv":
            mods[name] = nn.ConvTranspose2d(cin, cout, k, s, p, bias=True)
        else:
            mods[name] = nn.Conv2d(cin, cout, kernel_size=k, stride=s, padding=p, dilation=d, bias=True)
    for m in mods.values():
        nn.init.kaiming_normal_(m.weight.data, mode="fan_in")
        m.bias.data.zero_()
    sd = OrderedDict()
    for (name, kind, *_), m in zip(layer_specs(), mods.values()):
        key = param_key(name, kind)
        sd[f"{key}.weight"] = m.weight.data.clone()
        sd[f"{key}.bias"] = m.bias.data.clone()
    return sd


def randomize_biases(sd, seed=123, scale=0.05):
    """Tests use non-zero biases too (the reference init zeroes them, which would
    hide a bias bug)."""
    g = torch.Generator().manual_seed(seed)
    out = OrderedDict()
    for k, v in sd.items():
        out[k] = (torch.randn(v.shape, generator=g) * scale) if k.endswith(".bias") else v.clone()
    return out


# ----------------------------------------------------------------------------------------------
# operators
# ----------------------------------------------------------------------------------------------
def correlation_mean(f1, f2, md=MD):
    """81-channel cost volume with correlation_package (CUDA kernel) semantics.

    out[n,(dy+md)*(2md+1)+(dx+md),y,x] = (1/C) * sum_c f1[n,c,y,x] * f2[n,c,y+dy,x+dx], zero outside.
    Follows the shift-and-dot of correlation.py:22-39 and divides by ``nelems = k*k*C``
    as correlation_cuda_kernel.cu:104,143 does (corr_multiply is ignored by the
    CUDA kernel: correlation_cuda_kernel.cu:369).
    """
    B, C, H, W = f1.shape
    f2p = F.pad(f2, (md, md, md, md))
    outs = []
    for dy in range(-md, md + 1):
        for dx in range(-md, md + 1):
            patch = f2p[:, :, md + dy: md + dy + H, md + dx: md + dx + W]
            outs.append((f1 * patch).sum(dim=1, keepdim=True))
    return torch.cat(outs, dim=1) / float(C)


def correlation_backward_numpy(f1, f2, gout, md=MD):
    """(grad_f1, grad_f2) of correlation_mean, restated from correlation_cuda_kernel.cu:150-334 with numpy slices
    (fp64 accumulation):
      grad_f1[n,c,y,x] = (1/C) sum_{dy,dx} gout[n,tc,y,x]       * f2[n,c,y+dy,x+dx]      (:150-237, correlation_backward_input1)
      grad_f2[n,c,y,x] = (1/C) sum_{dy,dx} gout[n,tc,y-dy,x-dx] * f1[n,c,y-dy,x-dx]      (:241-334, correlation_backward_input2)
    tc = (dy+md)*(2md+1) + (dx+md); terms whose displaced pixel is outside the image vanish (zero padding)."""
    import numpy as np
    f1, f2, gout = (np.asarray(a, dtype=np.float64) for a in (f1, f2, gout))
    B, C, H, W = f1.shape
    g1, g2 = np.zeros_like(f1), np.zeros_like(f2)
    nd = 2 * md + 1
    for dy in range(-md, md + 1):
        for dx in range(-md, md + 1):
            t = (dy + md) * nd + dx + md
            ya, yb = max(0, -dy), min(H, H - dy)            # pixels (y,x) whose partner (y+dy, x+dx) is inside
            xa, xb = max(0, -dx), min(W, W - dx)
            if ya >= yb or xa >= xb:
                continue
            g = gout[:, t:t + 1, ya:yb, xa:xb]
            g1[:, :, ya:yb, xa:xb] += g * f2[:, :, ya + dy:yb + dy, xa + dx:xb + dx]
            g2[:, :, ya + dy:yb + dy, xa + dx:xb + dx] += g * f1[:, :, ya:yb, xa:xb]
    return (g1 / C).astype(np.float32), (g2 / C).astype(np.float32)


def warp(x, flo, align_corners=False, mask_thresh=0.9999):
    """PWCDCNet.warp (PWCNet.py:141-177) without the hard-wired ``.cuda()``.

    ``align_corners=False`` is what ``F.grid_sample`` defaults to under torch>=1.3, i.e. how
    the reference runs today; the grid is nevertheless normalised with the
    (W-1)/(H-1) formula, exactly as PWCNet.py:162-163 does.
    """
    B, C, H, W = x.shape
    xx = torch.arange(0, W).view(1, -1).repeat(H, 1)
    yy = torch.arange(0, H).view(-1, 1).repeat(1, W)
    xx = xx.view(1, 1, H, W).repeat(B, 1, 1, 1)
    yy = yy.view(1, 1, H, W).repeat(B, 1, 1, 1)
    grid = torch.cat((xx, yy), 1).float()
    vgrid = grid + flo
    vgrid[:, 0, :, :] = 2.0 * vgrid[:, 0, :, :].clone() / max(W - 1, 1) - 1.0
    vgrid[:, 1, :, :] = 2.0 * vgrid[:, 1, :, :].clone() / max(H - 1, 1) - 1.0
    vgrid = vgrid.permute(0, 2, 3, 1)
    output = F.grid_sample(x, vgrid, align_corners=align_corners)
    mask = F.grid_sample(torch.ones_like(x), vgrid, align_corners=align_corners)
    mask[mask < mask_thresh] = 0
    mask[mask > 0] = 1
    return output * mask


def warp_mask_sum(flo, H, W, align_corners=False):
    """The quantity PWCNet.py:168-174 thresholds at 0.9999: grid_sample(ones, vgrid) per pixel, [B,H,W] (test aid: a
    pixel whose sum sits within rounding of the threshold may legitimately flip between two fp32 evaluation orders)."""
    B = flo.shape[0]
    xx = torch.arange(0, W).view(1, -1).repeat(H, 1).view(1, 1, H, W).repeat(B, 1, 1, 1)
    yy = torch.arange(0, H).view(-1, 1).repeat(1, W).view(1, 1, H, W).repeat(B, 1, 1, 1)
    vgrid = torch.cat((xx, yy), 1).float() + flo
    vgrid[:, 0, :, :] = 2.0 * vgrid[:, 0, :, :].clone() / max(W - 1, 1) - 1.0
    vgrid[:, 1, :, :] = 2.0 * vgrid[:, 1, :, :].clone() / max(H - 1, 1) - 1.0
    return F.grid_sample(torch.ones(B, 1, H, W), vgrid.permute(0, 2, 3, 1), align_corners=align_corners)[:, 0]


def warp_corr_lrelu(f1, f2, flow, scale, align_corners=False, slope=0.1, mask_thresh=0.9999):
    """The fused op the CUDA kernel implements, composed from the reference pieces:
    leakyRELU(corr(f1, warp(f2, flow*scale)))  (PWCNet.py:212-214).  flow=None -> level 6 (PWCNet.py:198-199)."""
    w2 = f2 if flow is None else warp(f2, flow * scale, align_corners, mask_thresh)
    return F.leaky_relu(correlation_mean(f1, w2), slope)


def warp_corr_lrelu_numpy(f1, f2, flow, scale, align_corners=False, slope=0.1, mask_thresh=0.9999, md=MD):
    """Torch-free restatement of warp + mask + cost volume + LeakyReLU in fp32 numpy.

    Coordinates follow PWCNet.py:162-163 and ATen's grid_sampler_unnormalize
    (torch/include/ATen/native/GridSampler.h:27-35); corner weights and the zeros-padding
    accumulation follow ATen/native/cuda/GridSampler.cu (nw,ne,sw,se order).  SURVEY.md Appendix C.
    f1,f2: [B,C,H,W] float32; flow: [B,2,H,W] or None.
    """
    f1 = np.asarray(f1, np.float32); f2 = np.asarray(f2, np.float32)
    B, C, H, W = f1.shape
    f32 = np.float32
    if flow is None:
        w2 = f2
    else:
        flow = np.asarray(flow, np.float32)
        xs = np.arange(W, dtype=np.float32)[None, None, :]
        ys = np.arange(H, dtype=np.float32)[None, :, None]
        vx = xs + flow[:, 0] * f32(scale)
        vy = ys + flow[:, 1] * f32(scale)
        gx = f32(2.0) * vx / f32(max(W - 1, 1)) - f32(1.0)
        gy = f32(2.0) * vy / f32(max(H - 1, 1)) - f32(1.0)
        if align_corners:
            ix = (gx + f32(1)) / f32(2) * f32(W - 1)
            iy = (gy + f32(1)) / f32(2) * f32(H - 1)
        else:
            ix = ((gx + f32(1)) * f32(W) - f32(1)) / f32(2)
            iy = ((gy + f32(1)) * f32(H) - f32(1)) / f32(2)
        x0 = np.floor(ix); y0 = np.floor(iy); x1 = x0 + 1; y1 = y0 + 1
        wts = [((x1 - ix) * (y1 - iy), y0, x0), ((ix - x0) * (y1 - iy), y0, x1),
               ((x1 - ix) * (iy - y0), y1, x0), ((ix - x0) * (iy - y0), y1, x1)]
        val = np.zeros((B, C, H, W), np.float32)
        m = np.zeros((B, H, W), np.float32)
        bidx = np.arange(B)[:, None, None]
        for (wgt, yc, xc) in wts:
            inb = (xc >= 0) & (xc <= W - 1) & (yc >= 0) & (yc <= H - 1)
            yi = np.clip(yc, 0, H - 1).astype(np.int64); xi = np.clip(xc, 0, W - 1).astype(np.int64)
            wz = np.where(inb, wgt, f32(0)).astype(np.float32)
            g = f2[bidx, :, yi, xi]                      # [B,H,W,C]
            val += np.transpose(g, (0, 3, 1, 2)) * wz[:, None]
            m += wz
        keep = (m >= f32(mask_thresh)).astype(np.float32)
        w2 = val * keep[:, None]
    w2p = np.pad(w2, ((0, 0), (0, 0), (md, md), (md, md)))
    out = np.empty((B, (2 * md + 1) ** 2, H, W), np.float32)
    k = 0
    for dy in range(-md, md + 1):
        for dx in range(-md, md + 1):
            out[:, k] = (f1 * w2p[:, :, md + dy: md + dy + H, md + dx: md + dx + W]).sum(1, dtype=np.float32) / f32(C)
            k += 1
    return np.where(out > 0, out, out * f32(slope)).astype(np.float32)


# ----------------------------------------------------------------------------------------------
# full forward
# ----------------------------------------------------------------------------------------------
_BF16_SIM = [False]   # forward(..., bf16_storage=True): model the bf16-conv mode's storage roundings (test aid)


def _q(t):
    return t.to(_BF16_SIM[0]).float() if _BF16_SIM[0] else t


def _conv(sd, name, x, stride=1, pad=1, dil=1, act=True):
    key = f"{name}.0" if act else name
    w = sd[f"{key}.weight"]
    if name != "conv1a":          # conv1a runs on CUDA cores with fp32 weights in the bf16 mode too
        w = _q(w)
    y = F.conv2d(x, w, sd[f"{key}.bias"], stride=stride, padding=pad, dilation=dil)
    return _q(F.leaky_relu(y, 0.1)) if act else y      # predict_flow / dc_conv7 outputs stay fp32


def _deconv(sd, name, x):
    return F.conv_transpose2d(x, sd[f"{name}.weight"], sd[f"{name}.bias"], stride=2, padding=1)


@torch.no_grad()
def forward(sd, x, align_corners=False, return_intermediates=False, bf16_storage=False, storage=None):
    """PWCDCNet.forward in eval mode (PWCNet.py:180-273).  x: [B,6,H,W] fp32, H,W multiples of 64.
    Returns flow2 [B,2,H/4,W/4] (units of px/20, PWCNet.py:272-273); optionally a dict of per-level tensors.
    bf16_storage=True rounds weights and stored activations to bf16 where the bf16-conv mode of the CUDA path does
    (fp32 arithmetic everywhere): a test aid that separates precision effects from bugs; not reference behaviour.
    storage='bf16' | 'f16' selects the 16-bit format of that model (conv_mode 'bf16' / 'f16')."""
    _BF16_SIM[0] = {"bf16": torch.bfloat16, "f16": torch.float16, None: (torch.bfloat16 if bf16_storage else False)}[storage]
    try:
        return _forward(sd, x, align_corners, return_intermediates)
    finally:
        _BF16_SIM[0] = False


def _forward(sd, x, align_corners, return_intermediates):
    inter = {}
    im1, im2 = x[:, :3], x[:, 3:]

    def pyramid(im):
        feats = []
        c = im
        for lvl in range(1, 7):
            names = (f"conv{lvl}a", f"conv{lvl}aa", f"conv{lvl}b") if lvl < 6 else ("conv6aa", "conv6a", "conv6b")
            c = _conv(sd, names[0], c, stride=2)
            c = _conv(sd, names[1], c)
            c = _conv(sd, names[2], c)
            feats.append(c)
        return feats

    c1 = pyramid(im1)
    c2 = pyramid(im2)
    up_flow = up_feat = None
    flow = None
    xfeat = None
    for lvl in (6, 5, 4, 3, 2):
        f1, f2 = c1[lvl - 1], c2[lvl - 1]
        if lvl == 6:
            corr = _q(F.leaky_relu(correlation_mean(f1, f2), 0.1))
            xfeat = corr
        else:
            w2 = warp(f2, up_flow * FLOW_SCALE[lvl], align_corners)
            corr = _q(F.leaky_relu(correlation_mean(f1, w2), 0.1))
            xfeat = torch.cat((corr, f1, _q(up_flow), _q(up_feat)), 1)
            inter[f"warp{lvl}"] = w2
        inter[f"corr{lvl}"] = corr
        for i in range(5):
            xfeat = torch.cat((_conv(sd, f"conv{lvl}_{i}", xfeat), xfeat), 1)
        flow = _conv(sd, f"predict_flow{lvl}", xfeat, act=False)
        inter[f"flow{lvl}"] = flow
        if lvl > 2:
            up_flow = _deconv(sd, f"deconv{lvl}", flow)
            up_feat = _deconv(sd, f"upfeat{lvl}", xfeat)
            inter[f"up_flow{lvl}"] = up_flow
            inter[f"up_feat{lvl}"] = up_feat
    inter["x2"] = xfeat
    y = _conv(sd, "dc_conv1", xfeat, pad=1, dil=1)
    y = _conv(sd, "dc_conv2", y, pad=2, dil=2)
    y = _conv(sd, "dc_conv3", y, pad=4, dil=4)
    y = _conv(sd, "dc_conv4", y, pad=8, dil=8)
    y = _conv(sd, "dc_conv5", y, pad=16, dil=16)
    y = _conv(sd, "dc_conv6", y, pad=1, dil=1)
    flow2 = flow + _conv(sd, "dc_conv7", y, act=False)
    for i, c in enumerate(c1):
        inter[f"c1{i + 1}"] = c
    for i, c in enumerate(c2):
        inter[f"c2{i + 1}"] = c
    if return_intermediates:
        return flow2, inter
    return flow2


def synthetic_input(B, H, W, seed=1):
    """torch.manual_seed(seed); torch.rand(B,6,H,W) (uniform [0,1) = /255 range)."""
    g = torch.Generator().manual_seed(seed)
    return torch.rand(B, 6, H, W, generator=g)


def shifted_pair_input(B, H, W, seed=2, max_shift=6.0):
    """More realistic pair: im2 is im1 resampled by a smooth synthetic flow, so the decoder sees
    real correlations (seeds 2,3)."""
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(B, 3, H // 16, W // 16, generator=g)
    im1 = F.interpolate(low, size=(H, W), mode="bicubic", align_corners=False).clamp(0, 1)
    im1 = (im1 + 0.15 * torch.rand(B, 3, H, W, generator=g)).clamp(0, 1)
    fl = (torch.rand(B, 2, H // 32, W // 32, generator=g) - 0.5) * 2 * max_shift
    fl = F.interpolate(fl, size=(H, W), mode="bicubic", align_corners=False)
    ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    gx = 2 * (xs[None] + fl[:, 0]) / (W - 1) - 1
    gy = 2 * (ys[None] + fl[:, 1]) / (H - 1) - 1
    im2 = F.grid_sample(im1, torch.stack((gx, gy), -1), mode="bilinear", padding_mode="border", align_corners=True)
    return torch.cat((im1, im2), 1).contiguous()


def corr_microbench_inputs(level, B, H, W, C, with_flow=True):
    """SURVEY.md 8(d) microbench tensors: f = leaky_relu(randn,0.1) seeds 10+L / 20+L; flow = 2*randn seed 30+L."""
    g1 = torch.Generator().manual_seed(10 + level)
    g2 = torch.Generator().manual_seed(20 + level)
    g3 = torch.Generator().manual_seed(30 + level)
    f1 = F.leaky_relu(torch.randn(B, C, H, W, generator=g1), 0.1)
    f2 = F.leaky_relu(torch.randn(B, C, H, W, generator=g2), 0.1)
    flow = 2.0 * torch.randn(B, 2, H, W, generator=g3) if with_flow else None
    return f1, f2, flow
