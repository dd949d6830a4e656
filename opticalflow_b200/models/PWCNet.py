"""Drop-in for /root/reference/models/PWCNet.py :: PWCDCNet / pwc_dc_net (inference path).

Same class name, constructor (``md=4``), parameter names / shapes (the ``pwc_net.pth.tar`` state_dict layout:
128 tensors, PWCNet.py:52-132), same random init (PWCNet.py:134-138) and the same ``forward(x[B,6,H,W])``
contract (PWCNet.py:180-273: returns flow2 at 1/4 resolution in px/20; the tuple flow2..flow6 in training mode).
The forward itself runs as hand-written sm_100a kernels through the C ABI (opticalflow_b200/engine.py); the
nn.Conv2d modules below only hold the parameters.  There is no CPU or pure-torch fallback.
"""
import numpy as np
import torch
import torch.nn as nn

from .correlation_package.correlation import Correlation
from .. import engine

__all__ = ["pwc_dc_net", "PWCDCNet"]


def conv(in_planes, out_planes, kernel_size=3, stride=1, padding=1, dilation=1):
    return nn.Sequential(
        nn.Conv2d(in_planes, out_planes, kernel_size=kernel_size, stride=stride,
                  padding=padding, dilation=dilation, bias=True),
        nn.LeakyReLU(0.1))


def predict_flow(in_planes):
    return nn.Conv2d(in_planes, 2, kernel_size=3, stride=1, padding=1, bias=True)


def deconv(in_planes, out_planes, kernel_size=4, stride=2, padding=1):
    return nn.ConvTranspose2d(in_planes, out_planes, kernel_size, stride, padding, bias=True)


class PWCDCNet(nn.Module):
    """PWC-DC net (dilated context network + DenseNet connections), B200 execution.

    Extra keyword arguments (not in the reference): ``conv_mode`` 'fp32' (exact-fp32 CUDA-core convolutions),
    'bf16' (tcgen05 tensor-core convolutions, bf16 storage, fp32 accumulate) or 'f16' (the same kernels with IEEE-half
    storage: ~7x closer to the fp32 flow at the same speed); ``align_corners`` False reproduces the reference as
    it runs under torch>=1.3, True the 2018 semantics.
    """

    def __init__(self, md=4, conv_mode="fp32", align_corners=False):
        super().__init__()
        if md != 4:
            raise NotImplementedError("only md=4 is built (PWCNet.py:45,71)")
        ch = engine.PYR_CH
        for lvl in range(1, 7):
            a, aa, b = (f"conv{lvl}a", f"conv{lvl}aa", f"conv{lvl}b") if lvl < 6 else ("conv6aa", "conv6a", "conv6b")
            setattr(self, a, conv(ch[lvl - 1], ch[lvl], kernel_size=3, stride=2))
            setattr(self, aa, conv(ch[lvl], ch[lvl], kernel_size=3, stride=1))
            setattr(self, b, conv(ch[lvl], ch[lvl], kernel_size=3, stride=1))
        self.corr = Correlation(pad_size=md, kernel_size=1, max_displacement=md, stride1=1, stride2=1, corr_multiply=1)
        self.leakyRELU = nn.LeakyReLU(0.1)
        nd = (2 * md + 1) ** 2
        dd = np.cumsum([128, 128, 96, 64, 32])
        for lvl, extra in ((6, 0), (5, 128 + 4), (4, 96 + 4), (3, 64 + 4), (2, 32 + 4)):
            od = nd + extra
            setattr(self, f"conv{lvl}_0", conv(od, 128))
            setattr(self, f"conv{lvl}_1", conv(od + int(dd[0]), 128))
            setattr(self, f"conv{lvl}_2", conv(od + int(dd[1]), 96))
            setattr(self, f"conv{lvl}_3", conv(od + int(dd[2]), 64))
            setattr(self, f"conv{lvl}_4", conv(od + int(dd[3]), 32))
            setattr(self, f"predict_flow{lvl}", predict_flow(od + int(dd[4])))
            setattr(self, f"deconv{lvl}", deconv(2, 2, kernel_size=4, stride=2, padding=1))
            if lvl > 2:
                setattr(self, f"upfeat{lvl}", deconv(od + int(dd[4]), 2, kernel_size=4, stride=2, padding=1))
        od = nd + 32 + 4
        self.dc_conv1 = conv(od + int(dd[4]), 128, kernel_size=3, stride=1, padding=1, dilation=1)
        self.dc_conv2 = conv(128, 128, kernel_size=3, stride=1, padding=2, dilation=2)
        self.dc_conv3 = conv(128, 128, kernel_size=3, stride=1, padding=4, dilation=4)
        self.dc_conv4 = conv(128, 96, kernel_size=3, stride=1, padding=8, dilation=8)
        self.dc_conv5 = conv(96, 64, kernel_size=3, stride=1, padding=16, dilation=16)
        self.dc_conv6 = conv(64, 32, kernel_size=3, stride=1, padding=1, dilation=1)
        self.dc_conv7 = predict_flow(32)

        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
                nn.init.kaiming_normal_(m.weight.data, mode="fan_in")
                if m.bias is not None:
                    m.bias.data.zero_()

        self.conv_mode = conv_mode
        self.align_corners = align_corners
        self._plans = {}
        self._weights_epoch = 0
        self.verify_weights = False      # True: checksum the parameters every forward (catches edits through .data)

    # -- execution ---------------------------------------------------------------------------------
    def _plan(self, B, H, W, device, video=False, instance=0):
        """instance > 0: another plan of the same shape with its own activation buffers and CUDA graph, for callers that keep
        several forwards in flight on different streams (driver.FlowPipeline(inflight=2))."""
        key = (B, H, W, str(device), self.conv_mode, bool(self.align_corners), bool(video), int(instance))
        plan = self._plans.get(key)
        if plan is None:
            plan = engine.make_plan(self, B, H, W, device, self.conv_mode, self.align_corners, video=video)
            self._plans[key] = plan
        return plan

    def release_plans(self):
        self._plans.clear()

    def invalidate_weights(self):
        """Call after changing parameters through ``.data`` (``m.weight.data.copy_(...)``, ``nn.init.*(m.weight.data)``):
        such edits bump neither ``_version`` nor ``data_ptr``, so the packed kernel-side weights would go stale.
        ``load_state_dict`` calls this itself; ``verify_weights = True`` detects such edits automatically."""
        self._weights_epoch += 1

    def load_state_dict(self, *a, **kw):
        r = super().load_state_dict(*a, **kw)
        self.invalidate_weights()
        return r

    @torch.no_grad()
    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("PWCDCNet (B200): CUDA input required -- this build has no CPU fallback")
        if x.dim() != 4 or x.shape[1] != 6:
            raise RuntimeError("PWCDCNet expects x of shape [B,6,H,W] (PWCNet.py:181-182)")
        x = x.contiguous().float()
        B, _, H, W = x.shape
        with torch.cuda.device_of(x):
            plan = self._plan(B, H, W, x.device)
            plan.pack_weights(self)
            flow2 = plan.run(x).clone()
            if self.training:
                return (flow2,) + tuple(plan.level_flow(l) for l in (3, 4, 5, 6))
        return flow2

    def warp(self, x, flo):
        """PWCNet.py:141-177 as a standalone op on NCHW tensors (the forward uses the fused kernel instead):
        returns grid_sample(x, grid+flo) * mask, one kernel (pwc_warp_fwd) sharing the fused kernels' sampling code."""
        from ..ops import warp as _warp
        return _warp(x, flo, align_corners=self.align_corners)


def pwc_dc_net(path=None, **kw):
    """PWCNet.py:497-506: accepts {'state_dict': ...} or a flat state dict; also 'model' and a 'module.' prefix
    as the drivers do (pwc_extract_flow_video.py:167-172)."""
    model = PWCDCNet(**kw)
    if path is not None:
        data = torch.load(path, map_location="cpu")
        if isinstance(data, dict):
            data = data.get("state_dict") or data.get("model") or data
        data = {k.replace("module.", ""): v for k, v in data.items()}
        model.load_state_dict(data)
    return model
