"""Host-side helpers of the split-precision ('mixed') conv mode: an fp32-grade tensor is two 16-bit planes hi = RN16(v),
lo = RN16(v - hi); a convolution over it runs x_hi*W_hi + x_hi*W_lo + x_lo*W_hi on the tensor cores (include/pwc_b200.h,
"Split precision").  Pure tensor bookkeeping: which packed weight row multiplies which stored channel."""
import torch


def planes(t, dtype):
    """fp32 tensor -> (hi, lo) in `dtype` with hi + lo == t to ~2^-22 relative."""
    hi = t.to(dtype)
    lo = (t - hi.float()).to(dtype)
    return hi, lo


def residual(w, dtype):
    """W_lo as an fp32 tensor: the packer's RN16 of it is exact up to the second rounding."""
    return w - w.to(dtype).float()


def virtual_weights(w, dtype, kc, span, cmap=None, products=3):
    """w [rows, cin, 3, 3] fp32 whose input channel i sits at position cmap[i] (default i) of a `span`-channel slice that the
    kernel walks in chunks of kc.  Returns (w_virtual [rows, 3*cin, 3, 3], cmap_virtual, span_virtual) for
    pwc_pack_conv_weight_bf16: virtual chunk 3*ch multiplies x_hi by W_hi, 3*ch+1 x_hi by W_lo, 3*ch+2 x_lo by W_hi."""
    cin = w.shape[1]
    pos = torch.arange(cin) if cmap is None else torch.as_tensor(cmap, dtype=torch.long)
    ch, k = pos // kc, pos % kc
    nch = (span + kc - 1) // kc
    if products == 2:        # split activations, single-precision weights: 2*ch multiplies x_hi by W_hi, 2*ch+1 x_lo by W_hi
        return torch.cat([w, w], 1).contiguous(), torch.cat([(2 * ch) * kc + k, (2 * ch + 1) * kc + k]).tolist(), 2 * kc * nch
    wv = torch.cat([w, residual(w, dtype), w], 1).contiguous()
    cm = torch.cat([(3 * ch) * kc + k, (3 * ch + 1) * kc + k, (3 * ch + 2) * kc + k]).tolist()
    return wv, cm, 3 * kc * nch


def kpad_weights(w, dtype, idx, lo_pos, dup_pos, cmap):
    """K-padding form of the split for a SINGLE-precision consumer: the input channels `idx` (list) of w [rows, cin, 3, 3]
    are stored three times -- hi at cmap[i] (as before), lo at lo_pos[j], a second hi copy at dup_pos[j] -- and the layer
    gains 2*len(idx) input channels with weights W (rounded to W_hi by the packer) for the lo copies and W_lo for the hi
    copies.  Returns (w_ext, cmap_ext)."""
    wsel = w[:, idx]
    w_ext = torch.cat([w, wsel, residual(wsel, dtype)], 1).contiguous()
    return w_ext, list(cmap) + list(lo_pos) + list(dup_pos)
