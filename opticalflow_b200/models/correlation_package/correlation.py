"""Drop-in for /root/reference/models/correlation_package/correlation.py (Correlation, CorrelationFunction).

Same constructor arguments and call signature as correlation.py:92-117; the forward calls the sm_100a kernel
through the C ABI (``pwc_corr_fwd``, include/pwc_b200.h) instead of ``correlation_cuda.forward``
(correlation.py:70).  Inference only: there is no backward (the reference drivers run under no_grad:
inference.py:193, pwc_extract_flow_video.py:191).  There is no CPU or pure-torch fallback.
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


class CorrelationFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input1, input2, pad_size=3, kernel_size=3, max_displacement=20, stride1=1, stride2=2,
                corr_multiply=1):
        return correlation_forward(input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2,
                                   corr_multiply)

    @staticmethod
    def backward(ctx, grad_output):
        raise NotImplementedError("correlation backward is out of scope (inference path)")


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
