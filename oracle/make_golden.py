"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

    python -m oracle.make_golden

For each case: run /root/reference's PWCDCNet (via oracle/ref_harness.py) on a seeded synthetic
input with seeded random-init weights, record the final flow, per-level cost volumes / flows and
checksums of the weights and input.  tests/test_oracle.py re-creates weights and input from the
seeds on any box, checks the checksums, and compares oracle/pwcnet_oracle.py against these vectors.
"""
import hashlib
import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import pwcnet_oracle as O  # noqa: E402
from oracle import ref_harness  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = [
    # name, B, H, W, weight seed, input seed, input kind, random biases
    ("rand_64x128", 2, 64, 128, 0, 1, "rand", False),
    ("shift_128x192_bias", 1, 128, 192, 0, 2, "shift", True),
]


def digest(t):
    return hashlib.sha256(t.detach().contiguous().numpy().tobytes()).hexdigest()


def sd_digest(sd):
    h = hashlib.sha256()
    for k in sd:
        h.update(k.encode())
        h.update(sd[k].detach().contiguous().numpy().tobytes())
    return h.hexdigest()


def main():
    warnings.filterwarnings("ignore")
    os.makedirs(OUT, exist_ok=True)
    for (name, B, H, W, wseed, iseed, kind, rbias) in CASES:
        ref = ref_harness.reference_model(wseed)
        sd = ref.state_dict()
        if rbias:
            sd = O.randomize_biases(sd)
            ref.load_state_dict(sd)
        x = O.synthetic_input(B, H, W, iseed) if kind == "rand" else O.shifted_pair_input(B, H, W, iseed)
        grabbed = {}
        corr_calls = []
        h1 = ref.leakyRELU.register_forward_hook(lambda m, i, o: corr_calls.append(o.detach().clone()))
        for lvl in (2, 3, 4, 5, 6):
            getattr(ref, f"predict_flow{lvl}").register_forward_hook(
                lambda m, i, o, lvl=lvl: grabbed.__setitem__(f"flow{lvl}", o.detach().clone()))
        with torch.no_grad():
            flow2 = ref(x)
        h1.remove()
        # ref.leakyRELU is called once per level on the cost volume, level 6 first (PWCNet.py:199,214,...)
        for lvl, c in zip((6, 5, 4, 3, 2), corr_calls):
            grabbed[f"corr{lvl}"] = c
        # oracle state dict must be bit-identical to the reference init
        osd = O.reference_like_state_dict(wseed)
        if rbias:
            osd = O.randomize_biases(osd)
        assert list(osd.keys()) == list(sd.keys()), "state_dict key order/names differ"
        assert sd_digest(osd) == sd_digest(sd), "oracle init is not bit-identical to the reference init"
        out = {
            "flow2": flow2.numpy(),
            "meta": np.array([B, H, W, wseed, iseed, int(kind == "shift"), int(rbias)], np.int64),
            "sd_sha256": np.array(sd_digest(sd)),
            "x_sha256": np.array(digest(x)),
            "n_params": np.array(sum(v.numel() for v in sd.values())),
        }
        for lvl in (6, 5, 4):
            out[f"corr{lvl}"] = grabbed[f"corr{lvl}"].numpy()
        for lvl in (6, 5, 4, 3):
            out[f"flow{lvl}"] = grabbed[f"flow{lvl}"].numpy()
        # checksums only for the fine levels (keeps fixtures small)
        for lvl in (3, 2):
            c = grabbed[f"corr{lvl}"].double()
            out[f"corr{lvl}_stats"] = np.array([c.sum().item(), c.abs().sum().item(), c.abs().max().item()])
        path = os.path.join(OUT, f"{name}.npz")
        np.savez_compressed(path, **out)
        print(f"wrote {path}: flow2 {tuple(flow2.shape)} absmax {flow2.abs().max():.4f} "
              f"size {os.path.getsize(path)/1024:.1f} KiB")
    # standalone operator fixture: reference warp() + correlation on random tensors
    PWCDCNet, corr_mod = ref_harness.load_reference()
    torch.manual_seed(0)
    net = PWCDCNet().eval()
    f1, f2, flow = O.corr_microbench_inputs(4, 2, 12, 20, 24)
    with torch.no_grad():
        w2 = net.warp(f2, flow * 1.25)
        c = net.leakyRELU(net.corr(f1, w2))
        c6 = net.leakyRELU(net.corr(f1, f2))
    np.savez_compressed(os.path.join(OUT, "warp_corr_op.npz"), warp=w2.numpy(), corr=c.numpy(), corr_nowarp=c6.numpy(),
                        meta=np.array([4, 2, 12, 20, 24], np.int64))
    print("wrote warp_corr_op.npz")


if __name__ == "__main__":
    main()
