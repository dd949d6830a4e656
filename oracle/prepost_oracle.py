"""CPU restatement of the reference drivers' pre/post-processing, file formats and metrics.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and smoke()) as the checker of opticalflow_b200.prepost / flowio; never by
the product.  Pinned against outputs of the UNMODIFIED reference functions (extracted from /root/reference by
oracle/make_golden_prepost.py and run in this container) stored in tests/golden/prepost.npz -- see tests/test_prepost.py.
Each function cites the reference lines it follows.
"""
import numpy as np
import torch
import torch.nn.functional as F


def frames_to_input(frame1, frame2, swap_rb=True, multiple=64):
    """frame_to_tensor (pwc_extract_flow_video.py:27-34: BGR->RGB, astype(float32)/255, HWC->CHW), unsqueeze,
    pad_to_multiple_of_64 (:36-42, replicate, bottom/right), torch.cat([t1p, t2p], 1) (:204).  uint8 [H,W,3] inputs."""
    ts = []
    for fr in (frame1, frame2):
        a = np.asarray(fr)
        if swap_rb:
            a = a[:, :, ::-1]
        t = torch.from_numpy(np.transpose(a.astype(np.float32) / 255.0, (2, 0, 1)).copy())[None]
        ts.append(t)
    h, w = ts[0].shape[-2:]
    ph, pw = (multiple - h % multiple) % multiple, (multiple - w % multiple) % multiple
    ts = [F.pad(t, (0, pw, 0, ph), mode="replicate") for t in ts]
    return torch.cat(ts, 1), ph, pw


def frames_to_input_normalized(frame1, frame2, mean, std, swap_rb=False, multiple=64):
    """The KITTI drivers' transform (inference_kitti.py:175-179, inference.py:305-310): T.ToTensor -- HWC uint8 -> CHW,
    .float().div(255) -- then T.Normalize(mean, std) -- tensor.sub_(mean[:,None,None]).div_(std[:,None,None]) with fp32 mean / std
    (torchvision.transforms.functional.normalize) --, followed by the replicate pad + cat of model_infer (inference_kitti.py:53-63,
    216-217).  torchvision is not installed in this image: its two transforms are restated with the torch ops they are made of;
    PIL delivers RGB, hence swap_rb=False."""
    m = torch.as_tensor(mean, dtype=torch.float32)[:, None, None]
    sd = torch.as_tensor(std, dtype=torch.float32)[:, None, None]
    ts = []
    for fr in (frame1, frame2):
        a = np.asarray(fr)
        if swap_rb:
            a = a[:, :, ::-1]
        t = torch.from_numpy(np.ascontiguousarray(a)).permute(2, 0, 1).contiguous().float().div(255)
        ts.append(t.sub_(m).div_(sd)[None])
    h, w = ts[0].shape[-2:]
    ph, pw = (multiple - h % multiple) % multiple, (multiple - w % multiple) % multiple
    ts = [F.pad(t, (0, pw, 0, ph), mode="replicate") for t in ts]
    return torch.cat(ts, 1), ph, pw


def unpad(x, pad_h, pad_w):
    """inference_kitti.py:66-71."""
    if pad_h == 0 and pad_w == 0:
        return x
    H, W = x.shape[-2], x.shape[-1]
    return x[..., : H - pad_h, : W - pad_w]


def flow_resize(flow, new_h, new_w):
    """inference_kitti.py:80-91: bilinear, align_corners=True, u *= new_w/W, v *= new_h/H."""
    B, C, H, W = flow.shape
    if (H, W) == (new_h, new_w):
        return flow
    rs = F.interpolate(flow, size=(new_h, new_w), mode="bilinear", align_corners=True)
    rs[:, 0] *= new_w / W
    rs[:, 1] *= new_h / H
    return rs


def epe_metric(flow_pred, flow_gt, valid):
    """inference_kitti.py:94-107 ([H,W,2] float32 numpy, valid bool or None)."""
    du = flow_pred[..., 0] - flow_gt[..., 0]
    dv = flow_pred[..., 1] - flow_gt[..., 1]
    epe = np.sqrt(du * du + dv * dv)
    if valid is not None:
        epe = epe[valid]
    return float(np.mean(epe)) if epe.size else float("nan")


def fl_all_metric(flow_pred, flow_gt, valid):
    """inference_kitti.py:109-128 (EPE > max(3, 0.05 |gt|); == the AND form of inference.py:127-159)."""
    du = flow_pred[..., 0] - flow_gt[..., 0]
    dv = flow_pred[..., 1] - flow_gt[..., 1]
    epe = np.sqrt(du * du + dv * dv)
    mag = np.sqrt(flow_gt[..., 0] ** 2 + flow_gt[..., 1] ** 2)
    outlier = epe > np.maximum(3.0, 0.05 * mag)
    if valid is not None:
        outlier = outlier & valid
        denom = np.count_nonzero(valid)
    else:
        denom = outlier.size
    return 100.0 * float(np.count_nonzero(outlier)) / float(denom) if denom else float("nan")


def flo_bytes(uv):
    """script_pwc.py:12-27: tag, W, H, data."""
    uv = np.asarray(uv, dtype=np.float32)
    return (np.array(202021.25, np.float32).tobytes() + np.array(uv.shape[1], np.int32).tobytes() +
            np.array(uv.shape[0], np.int32).tobytes() + uv.tobytes())


def kitti_encode(flow):
    """inference.py:266-277: uint16 planes (valid, v, u) as saved -> returned here in R,G,B = (u, v, valid) order."""
    u = (flow[:, :, 0] * 64.0 + 2 ** 15).astype(np.uint16)
    v = (flow[:, :, 1] * 64.0 + 2 ** 15).astype(np.uint16)
    return np.stack([u, v, np.ones_like(u)], axis=-1)


def kitti_decode(rgb16):
    """inference_kitti.py:39-52 on an R,G,B uint16 array."""
    u = (rgb16[..., 0].astype(np.float32) - 32768.0) / 64.0
    v = (rgb16[..., 1].astype(np.float32) - 32768.0) / 64.0
    return np.stack([u, v], axis=-1).astype(np.float32), rgb16[..., 2] != 0
