"""Driver-side pre/post-processing and metrics on the GPU.

Same names / argument meaning as the reference drivers' helpers, but each is ONE CUDA kernel of libpwc_b200.so instead
of a chain of numpy / torch ops, and nothing here runs on the CPU (CUDA tensors required, no fallback):

  frames_to_input    frame_to_tensor + pad_to_multiple_of_64 + torch.cat   /root/reference/pwc_extract_flow_video.py:27-42,197-205
                                                                            /root/reference/inference_kitti.py:53-63,216-217
  unpad              /root/reference/inference_kitti.py:66-71, pwc_extract_flow_video.py:44-47 (same quirk: the pad sizes
                     are image pixels but are cut from the quarter-resolution flow)
  flow_resize        /root/reference/inference_kitti.py:80-91 (bilinear, align_corners=True, vectors rescaled)
  postprocess_flow   unpad + flow_resize (+ x20, script_pwc.py:72) fused
  epe_metric, fl_all_metric, flow_metrics   /root/reference/inference_kitti.py:94-128; inference.py:105-159
"""
import torch

from . import _lib


def pad_amounts(H, W, multiple=64):
    """(pad_h, pad_w) of pad_to_multiple_of_64 / pad_to_64 (bottom / right)."""
    return (multiple - H % multiple) % multiple, (multiple - W % multiple) % multiple


def _need_cuda(t, what):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise RuntimeError(f"{what}: CUDA tensors required (no CPU fallback)")


IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)     # inference_kitti.py:177-178


def frames_to_input(frames1, frames2, swap_rb=True, multiple=64, out=None, mean=None, std=None):
    """frames1, frames2: uint8 CUDA tensors [B,H,W,3] (or [H,W,3]) as cv2 delivers them (BGR).  Returns
    (x [B,6,Hp,Wp] fp32, pad_h, pad_w): x[:, :3] = frame 1 / 255 (RGB if swap_rb), x[:, 3:] = frame 2, replicate-padded
    to multiples of `multiple`.  mean / std (3 floats each, e.g. IMAGENET_MEAN / IMAGENET_STD): the KITTI drivers' variant,
    T.ToTensor followed by T.Normalize (inference_kitti.py:175-179; their PIL images are RGB: swap_rb=False)."""
    _need_cuda(frames1, "frames_to_input"); _need_cuda(frames2, "frames_to_input")
    if frames1.dim() == 3:
        frames1, frames2 = frames1[None], frames2[None]
    if frames1.dtype != torch.uint8 or frames1.shape != frames2.shape or frames1.shape[-1] != 3:
        raise RuntimeError("frames_to_input: two uint8 [B,H,W,3] tensors of the same shape expected")
    B, H, W, _ = frames1.shape
    dense = lambda t: t.stride()[1:] == (W * 3, 3, 1) and t.stride(0) >= H * W * 3      # noqa: E731  (batch stride is free)
    if not (dense(frames1) and dense(frames2) and frames1.stride(0) == frames2.stride(0)):
        frames1, frames2 = frames1.contiguous(), frames2.contiguous()
    img_stride = frames1.stride(0) if B > 1 else H * W * 3
    ph, pw = pad_amounts(H, W, multiple)
    Hp, Wp = H + ph, W + pw
    if Wp % 4:
        raise RuntimeError("frames_to_input: padded width must be a multiple of 4")
    if out is None:
        out = torch.empty(B, 6, Hp, Wp, device=frames1.device, dtype=torch.float32)
    lib = _lib.load()
    if (mean is None) != (std is None):
        raise RuntimeError("frames_to_input: mean and std come together")
    if mean is not None:
        import ctypes
        m3, s3 = (ctypes.c_float * 3)(*[float(v) for v in mean]), (ctypes.c_float * 3)(*[float(v) for v in std])
        e = lib.pwc_preprocess_u8_norm_fwd(_lib.ptr(frames1), _lib.ptr(frames2), img_stride, W * 3, B, H, W, Hp, Wp,
                                           int(bool(swap_rb)), m3, s3, _lib.ptr(out), _lib.stream_handle())
        _lib.check(e, "pwc_preprocess_u8_norm_fwd")
        return out, ph, pw
    e = lib.pwc_preprocess_u8_fwd(_lib.ptr(frames1), _lib.ptr(frames2), img_stride, W * 3, B, H, W, Hp, Wp,
                                  int(bool(swap_rb)), _lib.ptr(out), _lib.stream_handle())
    _lib.check(e, "pwc_preprocess_u8_fwd")
    return out, ph, pw


def frames_to_stack(frames, swap_rb=True, multiple=64, out=None):
    """frames: uint8 CUDA tensor [N,H,W,3] -> (x [N,3,Hp,Wp] fp32 = frame / 255, replicate-padded; pad_h, pad_w): the input
    of the shared-pyramid video plan (pair k = frames k and k+1)."""
    _need_cuda(frames, "frames_to_stack")
    if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3:
        raise RuntimeError("frames_to_stack: a uint8 [N,H,W,3] tensor expected")
    frames = frames.contiguous()
    N, H, W, _ = frames.shape
    ph, pw = pad_amounts(H, W, multiple)
    Hp, Wp = H + ph, W + pw
    if out is None:
        out = torch.empty(N, 3, Hp, Wp, device=frames.device, dtype=torch.float32)
    lib = _lib.load()
    e = lib.pwc_preprocess_frames_u8_fwd(_lib.ptr(frames), H * W * 3, W * 3, N, H, W, Hp, Wp, int(bool(swap_rb)), _lib.ptr(out),
                                         _lib.stream_handle())
    _lib.check(e, "pwc_preprocess_frames_u8_fwd")
    return out, ph, pw


def unpad(flow, pad_h, pad_w):
    """The reference's unpad (a view; no kernel)."""
    H, W = flow.shape[-2], flow.shape[-1]
    return flow[..., : H - pad_h, : W - pad_w]


def postprocess_flow(flow, pad_h=0, pad_w=0, size=None, mult=1.0, hwc=False):
    """flow [B,2,h,w] fp32 (CUDA) -> unpad(flow, pad_h, pad_w) resized to `size` = (H, W) (default: no resize), vectors
    rescaled by the size ratio and by `mult`; [B,2,H,W], or [B,H,W,2] with hwc=True."""
    _need_cuda(flow, "postprocess_flow")
    if flow.dim() != 4 or flow.shape[1] != 2 or flow.dtype != torch.float32:
        raise RuntimeError("postprocess_flow: fp32 [B,2,h,w] expected")
    flow = flow.contiguous()
    B, _, h, w = flow.shape
    ch, cw = h - pad_h, w - pad_w
    H, W = (ch, cw) if size is None else size
    out = torch.empty((B, H, W, 2) if hwc else (B, 2, H, W), device=flow.device, dtype=torch.float32)
    lib = _lib.load()
    e = lib.pwc_flow_postprocess_fwd(_lib.ptr(flow), B, h, w, ch, cw, _lib.ptr(out), H, W, float(mult), int(bool(hwc)),
                                     _lib.stream_handle())
    _lib.check(e, "pwc_flow_postprocess_fwd")
    return out


def flow_resize(flow, new_h, new_w):
    """Resize flow [B,2,H,W] to (new_h, new_w) and scale the vectors accordingly (inference_kitti.py:80-91)."""
    if tuple(flow.shape[-2:]) == (new_h, new_w):
        return flow
    return postprocess_flow(flow, 0, 0, (new_h, new_w))


def flow_metrics(flow_pred, flow_gt, valid=None, abs_thr=3.0, rel_thr=0.05):
    """Per-sample (EPE, Fl-all %, #valid) of [B,2,H,W] or [B,H,W,2] fp32 CUDA tensors; valid: [B,H,W] bool/uint8 or None.
    One device reduction; returns three CPU float64 tensors of length B (NaN where no pixel is valid, as
    inference_kitti.py:105-106,126-127)."""
    _need_cuda(flow_pred, "flow_metrics"); _need_cuda(flow_gt, "flow_metrics")
    if flow_pred.shape != flow_gt.shape or flow_pred.dim() != 4:
        raise RuntimeError("flow_metrics: prediction and ground truth must be 4-D tensors of the same shape")
    hwc = flow_pred.shape[-1] == 2 and flow_pred.shape[1] != 2
    B = flow_pred.shape[0]
    H, W = (flow_pred.shape[1], flow_pred.shape[2]) if hwc else (flow_pred.shape[2], flow_pred.shape[3])
    p, g = flow_pred.float().contiguous(), flow_gt.float().contiguous()
    v = None
    if valid is not None:
        _need_cuda(valid, "flow_metrics")
        v = (valid != 0).to(torch.uint8).contiguous()
    acc = torch.empty(B, 3, device=p.device, dtype=torch.float64)
    lib = _lib.load()
    e = lib.pwc_flow_metrics_fwd(_lib.ptr(p), _lib.ptr(g), _lib.ptr(v), B, H, W, int(hwc), float(abs_thr), float(rel_thr),
                                 _lib.ptr(acc), _lib.stream_handle())
    _lib.check(e, "pwc_flow_metrics_fwd")
    a = acc.cpu()
    n = a[:, 1]
    nan = torch.full_like(n, float("nan"))
    return torch.where(n > 0, a[:, 0] / n, nan), torch.where(n > 0, 100.0 * a[:, 2] / n, nan), n


def epe_metric(flow_pred, flow_gt, valid=None):
    """Average EPE over valid pixels of one sample ([H,W,2] or [2,H,W] CUDA tensors): inference_kitti.py:94-107."""
    return float(flow_metrics(flow_pred[None], flow_gt[None], None if valid is None else valid[None])[0][0])


def fl_all_metric(flow_pred, flow_gt, valid=None):
    """KITTI Fl-all outlier percentage of one sample: inference_kitti.py:109-128 (== inference.py:127-159)."""
    return float(flow_metrics(flow_pred[None], flow_gt[None], None if valid is None else valid[None])[1][0])
