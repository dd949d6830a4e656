"""Drop-in for /root/reference/models/correlation_package/correlation.py (Correlation, CorrelationFunction).

Same constructor arguments and call signature as correlation.py:92-117; the forward calls the sm_100a kernel
through the C ABI (``pwc_corr_fwd``, include/pwc_b200.h) instead of ``correlation_cuda.forward``
(correlation.py:70), the backward calls ``pwc_corr_bwd`` instead of ``correlation_cuda.backward`` (correlation.py:86-90).
There is no CPU or pure-torch fallback.
"""
import torch
import torch.nn as nn

from ... import _lib


def correlation_forward(input1, input2, pad_size=4, kernel_size=1, max_displacement=4, stride1=1, stride2=1,
                        corr_multiply=1):
    """out[n,(dy+md)*(2md+1)+(dx+md),y,x] = mean_c input1[n,c,y,x]*input2[n,c,y+dy,x+dx]
    (correlation_cuda_kernel.cu:73-147; corr_multiply is accepted and ignored exactly as the CUDA kernel does)."""
    if not (pad_size == max_displacement == 4 and kernel_size == 1 and stride1 == 1 and stride2 == 1):
        raise NotImplementedError("only the PWC-Net configuration pad=md=4, k=1, strides 1 is built (PWCNet.py:71)")
    if not (input1.is_cuda and input2.is_cuda):
        raise RuntimeError("Correlation: CUDA tensors required (no CPU fallback)")
    if input1.shape != input2.shape or input1.dim() != 4:
        raise RuntimeError("Correlation: inputs must be two [B,C,H,W] tensors of the same shape")
    if input1.dtype != torch.float32 or input2.dtype != torch.float32:
        raise RuntimeError("Correlation: float32 inputs required")
    lib = _lib.load()
    a, b = input1.contiguous(), input2.contiguous()
    B, C, H, W = a.shape
    out = torch.empty(B, (2 * max_displacement + 1) ** 2, H, W, device=a.device, dtype=a.dtype)
    with torch.cuda.device_of(a):
        err = lib.pwc_corr_fwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), B, C, H, W, max_displacement,
                               _lib.stream_handle())
    if err != 0:
        raise RuntimeError("CUDA call failed")   # same message as correlation_cuda.cc:81-83
    return out


def correlation_backward(input1, input2, grad_output, max_displacement=4, need1=True, need2=True):
    """(grad_input1, grad_input2) of correlation_forward (correlation_cuda_kernel.cu:150-334); None where not needed."""
    if not (input1.is_cuda and input2.is_cuda and grad_output.is_cuda):
        raise RuntimeError("Correlation: CUDA tensors required (no CPU fallback)")
    if max_displacement != 4:
        raise NotImplementedError("only the PWC-Net configuration pad=md=4, k=1, strides 1 is built (PWCNet.py:71)")
    B, C, H, W = input1.shape
    if tuple(grad_output.shape) != (B, (2 * max_displacement + 1) ** 2, H, W) or grad_output.dtype != torch.float32:
        raise RuntimeError("Correlation backward: grad_output must be float32 [B,81,H,W]")
    lib = _lib.load()
    a, b, g = input1.contiguous(), input2.contiguous(), grad_output.contiguous()
    g1 = torch.empty_like(a) if need1 else None
    g2 = torch.empty_like(b) if need2 else None
    with torch.cuda.device_of(a):
        err = lib.pwc_corr_bwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(g), _lib.ptr(g1), _lib.ptr(g2), B, C, H, W,
                               max_displacement, _lib.stream_handle())
    if err != 0:
        raise RuntimeError("CUDA call failed")   # correlation_cuda.cc:166-168
    return g1, g2


class CorrelationFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input1, input2, pad_size=3, kernel_size=3, max_displacement=20, stride1=1, stride2=2,
                corr_multiply=1):
        out = correlation_forward(input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2,
                                  corr_multiply)
        ctx.save_for_backward(input1, input2)
        ctx.max_displacement = max_displacement
        return out

    @staticmethod
    def backward(ctx, grad_output):
        input1, input2 = ctx.saved_tensors
        g1, g2 = correlation_backward(input1, input2, grad_output.float(), ctx.max_displacement,
                                      ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        return g1, g2, None, None, None, None, None, None


class Correlation(nn.Module):
    def __init__(self, pad_size, kernel_size, max_displacement, stride1, stride2, corr_multiply=1):
        super().__init__()
        self.pad_size = pad_size
        self.kernel_size = kernel_size
        self.max_displacement = max_displacement
        self.stride1 = stride1
        self.stride2 = stride2
        self.corr_multiply = corr_multiply

    def forward(self, input1, input2):
        return CorrelationFunction.apply(input1, input2, self.pad_size, self.kernel_size, self.max_displacement,
                                         self.stride1, self.stride2, self.corr_multiply)
