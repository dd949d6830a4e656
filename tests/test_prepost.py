"""pre/post-processing, file formats, metrics.
CPU: the oracle restatement and the host-side codecs against goldens produced by the UNMODIFIED reference helpers
(oracle/make_golden_prepost.py).  GPU: the CUDA kernels (through the C ABI) against the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import prepost_oracle as PO
from opticalflow_b200 import flowio

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "prepost.npz"))


# ------------------------------------------------------------------ CPU: oracle and codecs vs the reference's outputs
def test_oracle_preprocess_matches_reference():
    x, ph, pw = PO.frames_to_input(G["pre_fr1"], G["pre_fr2"])
    assert (ph, pw) == tuple(G["pre_pad"])
    assert np.array_equal(x.numpy(), G["pre_x"])                      # bit-exact
    ramp = np.arange(256, dtype=np.float32) / np.float32(255.0)
    assert np.array_equal(G["pre_ramp"][0, 0], ramp)


def test_oracle_postprocess_matches_reference():
    fl = torch.from_numpy(G["post_flow"])
    un = PO.unpad(fl, 5, 3)
    assert tuple(un.shape) == tuple(G["post_unpad_shape"]) == tuple(G["post_video_unpad_shape"])
    assert np.array_equal(PO.flow_resize(un.clone(), 37, 50).numpy(), G["post_resized"])
    assert np.array_equal(PO.flow_resize(fl.clone(), 64, 64).numpy(), G["post_resized_full"])


def test_oracle_metrics_match_reference():
    p, g, v = G["met_pred"], G["met_gt"], G["met_valid"]
    assert PO.epe_metric(p, g, v) == G["met_epe"][0] and PO.epe_metric(p, g, None) == G["met_epe"][1]
    assert PO.fl_all_metric(p, g, v) == G["met_fl"][0] and PO.fl_all_metric(p, g, None) == G["met_fl"][1]
    assert 0 < G["met_fl"][0] < 100


def test_flo_codec_is_byte_exact(tmp_path):
    p = tmp_path / "a.flo"
    flowio.write_flo(p, G["flo_uv"])
    assert p.read_bytes() == G["flo_bytes"].tobytes() == PO.flo_bytes(G["flo_uv"])
    flowio.writeFlowFile(p, G["flo_uv"])
    assert p.read_bytes() == G["flo_bytes"].tobytes()
    assert np.array_equal(flowio.read_flow(p), G["flo_read"])
    (tmp_path / "bad.flo").write_bytes(b"\x00" * 20)
    with pytest.raises(ValueError):
        flowio.read_flow(tmp_path / "bad.flo")
    with pytest.raises(ValueError):
        flowio.write_flo(p, np.zeros((3, 4, 3), np.float32))


def test_kitti_png_codec(tmp_path):
    # a file written by cv2 (what the reference reads): our reader == the reference's load_flow_kitti_png
    p = tmp_path / "k.png"
    p.write_bytes(G["kitti_png"].tobytes())
    assert np.array_equal(flowio.read_png16(p), G["kitti_rgb"])
    fl, va = flowio.read_kitti_png(p)
    assert np.array_equal(fl, G["kitti_flow"]) and np.array_equal(va, G["kitti_valid"])
    # encode: same quantisation as inference.py:266-277; own writer -> own reader and -> cv2 (if present)
    flow = (np.random.default_rng(3).standard_normal((11, 13, 2)) * 50).astype(np.float32)
    enc = flowio.encode_kitti(flow)
    assert np.array_equal(enc, PO.kitti_encode(flow))
    q = tmp_path / "w.png"
    flowio.write_kitti_png(q, flow)
    assert np.array_equal(flowio.read_png16(q), enc)
    dec, va = flowio.read_kitti_png(q)
    assert va.all() and np.abs(dec - flow).max() <= 1.0 / 64 and np.array_equal(dec, PO.kitti_decode(enc)[0])
    try:
        import cv2
    except ImportError:
        return
    assert np.array_equal(cv2.imread(str(q), cv2.IMREAD_UNCHANGED)[..., ::-1], enc)
    # every PNG row filter: files written by cv2 with different strategies decode identically
    big = np.random.default_rng(4).integers(0, 65536, (40, 33, 3), dtype=np.uint16)
    big[10:30] = (np.arange(33)[None, :, None] * 700 + np.arange(20)[:, None, None] * 300).astype(np.uint16)
    for strat in (cv2.IMWRITE_PNG_STRATEGY_DEFAULT, cv2.IMWRITE_PNG_STRATEGY_FILTERED, cv2.IMWRITE_PNG_STRATEGY_RLE):
        r = tmp_path / f"s{strat}.png"
        cv2.imwrite(str(r), big[..., ::-1].copy(), [cv2.IMWRITE_PNG_STRATEGY, strat])
        assert np.array_equal(flowio.read_png16(r), big)


def test_prepost_requires_cuda():
    from opticalflow_b200 import prepost
    with pytest.raises(RuntimeError):
        prepost.frames_to_input(torch.zeros(1, 8, 8, 3, dtype=torch.uint8), torch.zeros(1, 8, 8, 3, dtype=torch.uint8))
    with pytest.raises(RuntimeError):
        prepost.postprocess_flow(torch.zeros(1, 2, 4, 4))
    with pytest.raises(RuntimeError):
        prepost.flow_metrics(torch.zeros(1, 2, 4, 4), torch.zeros(1, 2, 4, 4))
    assert prepost.pad_amounts(375, 1242) == (9, 38) and prepost.pad_amounts(1080, 1920) == (8, 0)


# ------------------------------------------------------------------ GPU: kernels vs the oracle / the goldens
@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 37, 50), (2, 64, 128), (3, 375, 1242), (1, 1080, 1920), (2, 5, 4)])
@pytest.mark.parametrize("swap", [True, False])
def test_gpu_preprocess_bit_exact(shape, swap):
    from opticalflow_b200 import prepost
    B, H, W = shape
    g = np.random.default_rng(B * 1000 + H)
    f1 = g.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    f2 = g.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    if shape == (1, 37, 50):
        f1[0], f2[0] = G["pre_fr1"], G["pre_fr2"]
    x, ph, pw = prepost.frames_to_input(torch.from_numpy(f1).cuda(), torch.from_numpy(f2).cuda(), swap_rb=swap)
    x = x.cpu()
    for n in range(B):
        ref, rph, rpw = PO.frames_to_input(f1[n], f2[n], swap_rb=swap)
        assert (ph, pw) == (rph, rpw)
        assert torch.equal(x[n], ref[0])                              # bit-exact (integer -> correctly rounded /255)
    if shape == (1, 37, 50) and swap:
        assert np.array_equal(x.numpy(), G["pre_x"])                  # the unmodified reference's output


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 37, 50), (2, 375, 1242)])
@pytest.mark.parametrize("swap", [False, True])
def test_gpu_preprocess_imagenet_normalised_bit_exact(shape, swap):
    """T.ToTensor + T.Normalize(ImageNet) of the KITTI drivers (inference_kitti.py:175-179) fused into the pre-processing kernel"""
    from opticalflow_b200 import prepost
    B, H, W = shape
    g = np.random.default_rng(77 + H)
    f1 = g.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    f2 = g.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    f1[0, 0, :4] = [[0, 0, 0], [255, 255, 255], [1, 128, 254], [124, 116, 104]]
    x, ph, pw = prepost.frames_to_input(torch.from_numpy(f1).cuda(), torch.from_numpy(f2).cuda(), swap_rb=swap,
                                        mean=prepost.IMAGENET_MEAN, std=prepost.IMAGENET_STD)
    x = x.cpu()
    for n in range(B):
        ref, rph, rpw = PO.frames_to_input_normalized(f1[n], f2[n], prepost.IMAGENET_MEAN, prepost.IMAGENET_STD, swap_rb=swap)
        assert (ph, pw) == (rph, rpw)
        assert torch.equal(x[n], ref[0])


@pytest.mark.gpu
@pytest.mark.parametrize("case", [
    # B, h, w, pad_h, pad_w, H, W, mult
    (2, 16, 16, 5, 3, 37, 50, 1.0), (2, 16, 16, 0, 0, 64, 64, 1.0), (1, 96, 320, 9, 38, 375, 1242, 20.0),
    (1, 7, 9, 0, 0, 7, 9, 20.0), (3, 5, 6, 1, 2, 1, 1, 1.0), (1, 272, 480, 8, 0, 1080, 1920, 20.0),
])
def test_gpu_postprocess_matches_oracle(case):
    from opticalflow_b200 import prepost
    B, h, w, ph, pw, H, W, mult = case
    g = torch.Generator().manual_seed(h * w)
    fl = torch.randn(B, 2, h, w, generator=g) * 3
    if case[:3] == (2, 16, 16):
        fl = torch.from_numpy(G["post_flow"])
    ref = PO.flow_resize((PO.unpad(fl, ph, pw) * mult).clone(), H, W)
    out = prepost.postprocess_flow(fl.cuda(), ph, pw, (H, W), mult)
    hwc = prepost.postprocess_flow(fl.cuda(), ph, pw, (H, W), mult, hwc=True)
    assert torch.equal(hwc.permute(0, 3, 1, 2), out)
    # fp32 bilinear blend: same formula as ATen, summation order / FMA contraction may differ -> 1e-5 of the flow scale
    tol = 1e-5 * max(1.0, ref.abs().max().item())
    assert (out.cpu() - ref).abs().max().item() <= tol
    if case == (2, 16, 16, 5, 3, 37, 50, 1.0):
        assert np.abs(out.cpu().numpy() - G["post_resized"]).max() <= tol
    assert prepost.flow_resize(fl.cuda(), h, w).data_ptr() is not None


@pytest.mark.gpu
def test_gpu_metrics_match_reference():
    from opticalflow_b200 import prepost
    p, g, v = G["met_pred"], G["met_gt"], G["met_valid"]
    tp, tg, tv = torch.from_numpy(p).cuda(), torch.from_numpy(g).cuda(), torch.from_numpy(v).cuda()
    assert abs(prepost.epe_metric(tp, tg, tv) - G["met_epe"][0]) <= 1e-6 * G["met_epe"][0]     # fp32 mean vs fp64 sum
    assert abs(prepost.epe_metric(tp, tg, None) - G["met_epe"][1]) <= 1e-6 * G["met_epe"][1]
    assert prepost.fl_all_metric(tp, tg, tv) == pytest.approx(G["met_fl"][0], rel=1e-12)      # integer counts: exact
    assert prepost.fl_all_metric(tp, tg, None) == pytest.approx(G["met_fl"][1], rel=1e-12)
    # batched NCHW at KITTI size, per-sample results, one empty validity mask
    B, H, W = 3, 375, 1242
    gen = torch.Generator().manual_seed(5)
    pred = torch.randn(B, 2, H, W, generator=gen) * 5
    gt = torch.randn(B, 2, H, W, generator=gen) * 5
    gt[:, :, :100] *= 40
    valid = torch.rand(B, H, W, generator=gen) > 0.5
    valid[2] = False
    epe, fl, n = prepost.flow_metrics(pred.cuda(), gt.cuda(), valid.cuda())
    for b in range(2):
        pn, gn = pred[b].permute(1, 2, 0).numpy(), gt[b].permute(1, 2, 0).numpy()
        assert n[b].item() == valid[b].sum().item()
        assert abs(epe[b].item() - PO.epe_metric(pn, gn, valid[b].numpy())) <= 1e-5 * epe[b].item()
        assert fl[b].item() == pytest.approx(PO.fl_all_metric(pn, gn, valid[b].numpy()), rel=1e-12)
    assert n[2].item() == 0 and np.isnan(epe[2].item()) and np.isnan(fl[2].item())
