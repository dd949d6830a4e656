"""Host-side weight transforms of the bf16 plan, checked on CPU against plain torch ops (no device code runs)."""
import pytest
import torch
import torch.nn.functional as F

from opticalflow_b200 import engine_bf16 as E


@pytest.mark.parametrize("g,C", [(4, 16), (2, 32)])
def test_space_to_depth_weights_reproduce_the_conv(g, C):
    torch.manual_seed(0)
    w, b = torch.randn(C, C, 3, 3, dtype=torch.float64), torch.randn(C, dtype=torch.float64)
    x = torch.randn(2, C, 6, 4 * g, dtype=torch.float64)
    want = F.conv2d(x, w, b, padding=1)
    w2, b2 = E._s2d_weights("k", g)({"k.weight": w, "k.bias": b})
    w2, b2 = w2.double(), b2.double()
    assert w2.shape == (g * C, g * C, 3, 3)
    N, _, H, W = x.shape
    # NHWC [N,H,W,C] viewed as [N,H,W/g,g*C] -> NCHW of the view: channel index = dx*C + c
    xs = x.permute(0, 2, 3, 1).reshape(N, H, W // g, g * C).permute(0, 3, 1, 2)
    ys = F.conv2d(xs, w2, b2, padding=1)
    got = ys.permute(0, 2, 3, 1).reshape(N, H, W, C).permute(0, 3, 1, 2)
    assert torch.allclose(got, want, atol=1e-5)          # w2 is built in fp32
    # density of the block matrix: each output uses 3 of the 3*g x-positions
    assert (w2 != 0).double().mean().item() == pytest.approx(3.0 / (3 * g), rel=1e-6)


@pytest.mark.parametrize("lvl", [3, 6])
def test_predict_flow_plus_upfeat_as_one_conv(lvl):
    torch.manual_seed(1)
    cin = 12
    sd = {f"predict_flow{lvl}.weight": torch.randn(2, cin, 3, 3), f"predict_flow{lvl}.bias": torch.randn(2),
          f"upfeat{lvl}.weight": torch.randn(cin, 2, 4, 4), f"upfeat{lvl}.bias": torch.randn(2)}
    x = torch.randn(1, cin, 5, 7)
    w, b = E._pfup_weights(lvl)(sd)
    y = F.conv2d(x, w, b, padding=1)
    assert torch.allclose(y[:, :2], F.conv2d(x, sd[f"predict_flow{lvl}.weight"], sd[f"predict_flow{lvl}.bias"], padding=1),
                          atol=1e-5)
    up = F.conv_transpose2d(x, sd[f"upfeat{lvl}.weight"], sd[f"upfeat{lvl}.bias"], stride=2, padding=1)
    for a in range(2):
        for bb in range(2):
            c0 = 2 + 2 * (2 * a + bb)
            assert torch.allclose(y[:, c0:c0 + 2], up[:, :, a::2, bb::2], atol=1e-4)


def test_pwc_dc_net_loader_accepts_the_checkpoint_layouts(tmp_path):
    """pwc_dc_net(path) (PWCNet.py:497-506) and the drivers' variants: {'state_dict': ...}, {'model': ...}, a flat dict,
    and keys carrying DataParallel's 'module.' prefix (pwc_extract_flow_video.py:167-172)."""
    import torch
    from opticalflow_b200.models.PWCNet import PWCDCNet, pwc_dc_net
    torch.manual_seed(3)
    src = PWCDCNet()
    sd = {k: v.clone() for k, v in src.state_dict().items()}
    assert len(sd) == 128
    variants = {
        "flat.pth.tar": sd,
        "state_dict.pth.tar": {"state_dict": sd, "epoch": 7},
        "model.pth.tar": {"model": sd},
        "module.pth.tar": {"state_dict": {"module." + k: v for k, v in sd.items()}},
    }
    for name, obj in variants.items():
        p = tmp_path / name
        torch.save(obj, p)
        m = pwc_dc_net(str(p))
        got = m.state_dict()
        assert set(got) == set(sd), name
        assert all(torch.equal(got[k], sd[k]) for k in sd), name
        assert m._weights_epoch >= 1                       # load_state_dict invalidates the packed kernel-side weights
    # a checkpoint of another topology is refused, as torch's strict load_state_dict does in the reference
    bad = dict(sd); bad.pop("conv1a.0.weight")
    torch.save(bad, tmp_path / "bad.pth.tar")
    import pytest
    with pytest.raises(RuntimeError):
        pwc_dc_net(str(tmp_path / "bad.pth.tar"))


def test_weights_key_sees_data_edits_when_asked():
    import torch
    from opticalflow_b200.engine import weights_key
    from opticalflow_b200.models.PWCNet import PWCDCNet
    m = PWCDCNet()
    k0 = weights_key(m)
    m.conv1a[0].weight.data.mul_(2.0)                      # the reference's init idiom: invisible to _version / data_ptr
    assert weights_key(m) == k0
    m.invalidate_weights()
    assert weights_key(m) != k0
    m.verify_weights = True
    k1 = weights_key(m)
    m.conv1a[0].weight.data.mul_(0.5)
    assert weights_key(m) != k1                            # the checksum catches it without an explicit call
