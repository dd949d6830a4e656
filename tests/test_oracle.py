"""CPU tests: the oracle (oracle/pwcnet_oracle.py) against the golden vectors produced from the
UNMODIFIED reference by oracle/make_golden.py, and its two restatements against each other."""
import hashlib
import os
import warnings

import numpy as np
import pytest
import torch

from oracle import pwcnet_oracle as O

warnings.filterwarnings("ignore")


def _sd_digest(sd):
    h = hashlib.sha256()
    for k in sd:
        h.update(k.encode())
        h.update(sd[k].detach().contiguous().numpy().tobytes())
    return h.hexdigest()


def _case(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    B, H, W, wseed, iseed, shift, rbias = [int(v) for v in g["meta"]]
    sd = O.reference_like_state_dict(wseed)
    if rbias:
        sd = O.randomize_biases(sd)
    x = O.shifted_pair_input(B, H, W, iseed) if shift else O.synthetic_input(B, H, W, iseed)
    return g, sd, x


@pytest.mark.parametrize("name", ["rand_64x128", "shift_128x192_bias"])
def test_oracle_matches_reference_golden(golden_dir, name):
    g, sd, x = _case(golden_dir, name)
    # weights and input are regenerated from seeds: they must be the very tensors the reference saw
    assert _sd_digest(sd) == str(g["sd_sha256"]), "random-init weights differ from the reference's"
    assert hashlib.sha256(x.numpy().tobytes()).hexdigest() == str(g["x_sha256"])
    assert sum(v.numel() for v in sd.values()) == int(g["n_params"]) == 9374340
    assert len(sd) == 128
    flow2, inter = O.forward(sd, x, return_intermediates=True)
    # same torch CPU kernels, same op order -> expect (near) bit equality; tolerance 1e-6 absolute
    np.testing.assert_allclose(flow2.numpy(), g["flow2"], rtol=0, atol=1e-6)
    for lvl in (6, 5, 4):
        np.testing.assert_allclose(inter[f"corr{lvl}"].numpy(), g[f"corr{lvl}"], rtol=0, atol=1e-6)
    for lvl in (6, 5, 4, 3):
        np.testing.assert_allclose(inter[f"flow{lvl}"].numpy(), g[f"flow{lvl}"], rtol=0, atol=1e-6)
    for lvl in (3, 2):
        c = inter[f"corr{lvl}"].double()
        st = np.array([c.sum().item(), c.abs().sum().item(), c.abs().max().item()])
        np.testing.assert_allclose(st, g[f"corr{lvl}_stats"], rtol=1e-6)


def test_oracle_ops_match_reference_warp_and_corr(golden_dir):
    g = np.load(os.path.join(golden_dir, "warp_corr_op.npz"))
    lvl, B, H, W, C = [int(v) for v in g["meta"]]
    f1, f2, flow = O.corr_microbench_inputs(lvl, B, H, W, C)
    w2 = O.warp(f2, flow * 1.25)
    np.testing.assert_allclose(w2.numpy(), g["warp"], rtol=0, atol=1e-7)
    c = O.warp_corr_lrelu(f1, f2, flow, 1.25)
    np.testing.assert_allclose(c.numpy(), g["corr"], rtol=0, atol=1e-6)
    c6 = O.warp_corr_lrelu(f1, f2, None, 1.0)
    np.testing.assert_allclose(c6.numpy(), g["corr_nowarp"], rtol=0, atol=1e-6)


def test_numpy_restatement_matches_reference(golden_dir):
    """The torch-free restatement (the formulas the CUDA kernel implements)
    reproduces the reference's warp+corr output; mask flips at the 0.9999 threshold are counted."""
    g = np.load(os.path.join(golden_dir, "warp_corr_op.npz"))
    lvl, B, H, W, C = [int(v) for v in g["meta"]]
    f1, f2, flow = O.corr_microbench_inputs(lvl, B, H, W, C)
    out = O.warp_corr_lrelu_numpy(f1.numpy(), f2.numpy(), flow.numpy(), 1.25)
    ref = g["corr"]
    scale = np.abs(ref).max()
    assert np.abs(out - ref).max() <= 1e-5 * scale
    out6 = O.warp_corr_lrelu_numpy(f1.numpy(), f2.numpy(), None, 1.0)
    assert np.abs(out6 - g["corr_nowarp"]).max() <= 1e-5 * np.abs(g["corr_nowarp"]).max()


def test_zero_flow_masks_border():
    """with zero flow and align_corners=False the reference zeroes a 1-px border."""
    f2 = torch.ones(1, 3, 8, 10)
    w = O.warp(f2, torch.zeros(1, 2, 8, 10))
    assert w[0, :, 0, :].abs().max() == 0 and w[0, :, -1, :].abs().max() == 0
    assert w[0, :, :, 0].abs().max() == 0 and w[0, :, :, -1].abs().max() == 0
    assert torch.allclose(w[0, :, 1:-1, 1:-1], torch.ones(3, 6, 8), atol=1e-6)
    w_ac = O.warp(f2, torch.zeros(1, 2, 8, 10), align_corners=True)
    assert torch.allclose(w_ac, f2, atol=1e-6)


def test_correlation_channel_order_and_mean():
    """channel = (dy+4)*9+(dx+4), displacement applied to input2, mean over C."""
    f1 = torch.zeros(1, 2, 12, 12); f2 = torch.zeros(1, 2, 12, 12)
    f1[0, :, 5, 5] = 1.0
    f2[0, :, 7, 4] = 3.0   # dy=+2, dx=-1
    c = O.correlation_mean(f1, f2)
    k = (2 + 4) * 9 + (-1 + 4)
    assert c[0, k, 5, 5].item() == pytest.approx(3.0)   # (1*3 + 1*3)/2
    c[0, k, 5, 5] = 0
    assert c.abs().max() == 0
