"""CPU error budget, part 4: which PYRAMID levels could drop to single half precision (weights and stored activations rounded)
with everything else as the mixed plan has it (B': level 5/4 decoders and up_flow/up_feat exact, the rest half)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import pwcnet_oracle as O

H, W = 448, 1024
torch.set_num_threads(8)
sd = O.reference_like_state_dict(0)
hf = torch.float16
q = lambda t: t.to(hf).float()


def fwd(x, corr_in_q=True, split_levels=(6, 5, 4), corr_out_q=True, dec_q=True, heads_q=True, c1_q=True, split_bits=None, pyr_q=()):
    """pyramid exact; optional roundings elsewhere. split_levels: decoder levels kept exact."""
    def sq(t):      # precision of a hi/lo split tensor (exact by default; split_bits emulates hi+lo with that many mantissa bits)
        return t
    def conv(name, t, stride=1, pad=1, dil=1, act=True, quant=False):
        key = f"{name}.0" if act else name
        w = sd[f"{key}.weight"]
        if quant: w = q(w)
        y = F.conv2d(t, w, sd[f"{key}.bias"], stride=stride, padding=pad, dilation=dil)
        if act: y = F.leaky_relu(y, 0.1)
        return q(y) if (quant and act) else y
    im1, im2 = x[:, :3], x[:, 3:]
    def pyr(im):
        fs = []; c = im
        for l in range(1, 7):
            n = (f"conv{l}a", f"conv{l}aa", f"conv{l}b") if l < 6 else ("conv6aa", "conv6a", "conv6b")
            ql = l in pyr_q
            c = conv(n[0], c, stride=2, quant=ql); c = conv(n[1], c, quant=ql); c = conv(n[2], c, quant=ql); fs.append(c)
        return fs
    c1, c2 = pyr(im1), pyr(im2)
    up_flow = up_feat = None
    for l in (6, 5, 4, 3, 2):
        f1, f2 = c1[l - 1], c2[l - 1]
        exact = l in split_levels
        a, b = (q(f1), q(f2)) if corr_in_q else (f1, f2)
        if l == 6:
            corr = F.leaky_relu(O.correlation_mean(a, b), 0.1)
        else:
            corr = F.leaky_relu(O.correlation_mean(a, O.warp(b, up_flow * O.FLOW_SCALE[l])), 0.1)
        if corr_out_q: corr = q(corr)
        f1d = q(f1) if (c1_q) else f1
        if l == 6:
            xf = corr
        else:
            uf, ufe = (q(up_flow), q(up_feat)) if heads_q else (up_flow, up_feat)
            xf = torch.cat((corr, f1d, uf, ufe), 1)
        for i in range(5):
            xf = torch.cat((conv(f"conv{l}_{i}", xf, quant=(dec_q and not exact)), xf), 1)
        wq = (dec_q and not exact)
        flow = F.conv2d(xf, q(sd[f"predict_flow{l}.weight"]) if wq else sd[f"predict_flow{l}.weight"], sd[f"predict_flow{l}.bias"], padding=1)
        if l > 2:
            up_flow = F.conv_transpose2d(flow, sd[f"deconv{l}.weight"], sd[f"deconv{l}.bias"], stride=2, padding=1)
            wu = sd[f"upfeat{l}.weight"]
            up_feat = F.conv_transpose2d(xf, q(wu) if wq else wu, sd[f"upfeat{l}.bias"], stride=2, padding=1)
    y = xf
    for n, d in (("dc_conv1", 1), ("dc_conv2", 2), ("dc_conv3", 4), ("dc_conv4", 8), ("dc_conv5", 16), ("dc_conv6", 1)):
        y = conv(n, y, pad=d, dil=d, quant=dec_q)
    return flow + F.conv2d(y, q(sd["dc_conv7.weight"]) if dec_q else sd["dc_conv7.weight"], sd["dc_conv7.bias"], padding=1)


def epe(a, b): return (a - b).pow(2).sum(1).sqrt().mean().item()

for name, x in (("rand", O.synthetic_input(1, H, W, 1)), ("shift", O.shifted_pair_input(1, H, W, 3))):
    ref = O.forward(sd, x)
    print(f"[{name}]")
    base = dict(heads_q=False, split_levels=(5, 4))
    for label, kw in {
        "mixed plan (pyramid exact)": dict(base),
        "pyramid level 1 half": dict(base, pyr_q=(1,)),
        "pyramid level 2 half": dict(base, pyr_q=(2,)),
        "pyramid levels 1-2 half": dict(base, pyr_q=(1, 2)),
        "pyramid level 6 half": dict(base, pyr_q=(6,)),
        "pyramid levels 5-6 half": dict(base, pyr_q=(5, 6)),
        "pyramid level 3 half": dict(base, pyr_q=(3,)),
    }.items():
        with torch.no_grad():
            out = fwd(x, **kw)
        print(f"  {label:62s} {20 * epe(out, ref):.4f} px", flush=True)
