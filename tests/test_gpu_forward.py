"""GPU parity of the whole PWCDCNet forward (fp32 mode) against the golden vectors of the UNMODIFIED reference and
against the CPU oracle on larger inputs.  Bar: final flow within 1e-3 px max-abs (flow2 is in px/20: x20)."""
import os
import warnings

import numpy as np
import pytest
import torch

from oracle import pwcnet_oracle as O

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu


def _model(wseed, rbias, **kw):
    from opticalflow_b200.models.PWCNet import PWCDCNet
    torch.manual_seed(wseed)
    m = PWCDCNet(**kw).eval()
    if rbias:
        m.load_state_dict(O.randomize_biases(m.state_dict()))
    return m.cuda()


@pytest.mark.parametrize("name", ["rand_64x128", "shift_128x192_bias"])
def test_forward_fp32_matches_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    B, H, W, wseed, iseed, shift, rbias = [int(v) for v in g["meta"]]
    m = _model(wseed, rbias)
    x = O.shifted_pair_input(B, H, W, iseed) if shift else O.synthetic_input(B, H, W, iseed)
    flow2 = m(x.cuda()).cpu()
    ref = torch.from_numpy(g["flow2"])
    assert flow2.shape == ref.shape
    err_px = 20.0 * (flow2 - ref).abs().max().item()
    assert err_px <= 1e-3, f"final flow differs by {err_px:.3e} px"
    plan = next(iter(m._plans.values()))
    for lvl in (6, 5, 4, 3):
        e = 20.0 * (plan.level_flow(lvl).cpu() - torch.from_numpy(g[f"flow{lvl}"])).abs().max().item()
        assert e <= 1e-3, f"flow{lvl} differs by {e:.3e} px"
    # cost volumes straight out of the concat buffer (channels 448..528)
    for lvl in (6, 5, 4):
        c = plan.cat[lvl][..., 448:529].permute(0, 3, 1, 2).cpu()
        refc = torch.from_numpy(g[f"corr{lvl}"])
        assert (c - refc).abs().max().item() <= 1e-4 * refc.abs().max().item()


@pytest.mark.parametrize("shape", [(1, 448, 1024), (2, 384, 1280)])
def test_forward_fp32_matches_oracle_full_size(shape):
    B, H, W = shape
    m = _model(0, True)
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    x = O.shifted_pair_input(B, H, W, 3)
    ref = O.forward(sd, x)
    flow2 = m(x.cuda()).cpu()
    err_px = 20.0 * (flow2 - ref).abs().max().item()
    assert err_px <= 1e-3, f"final flow differs by {err_px:.3e} px (flow absmax {20*ref.abs().max().item():.2f} px)"
    # second call reuses the plan and must be bit-identical (no stale state in the buffers)
    assert torch.equal(m(x.cuda()).cpu(), flow2)


def test_training_mode_returns_five_flows():
    m = _model(0, False).train()
    out = m(O.synthetic_input(1, 64, 64, 1).cuda())
    assert isinstance(out, tuple) and len(out) == 5
    assert [tuple(o.shape[2:]) for o in out] == [(16, 16), (8, 8), (4, 4), (2, 2), (1, 1)]
