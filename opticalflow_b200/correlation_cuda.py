"""Drop-in for the reference's pybind11 extension module ``correlation_cuda`` (models/correlation_package/correlation_cuda.cc:
10-16, 89-96, 169-172): same module-level functions, same argument lists, same in-place ownership contract, served by
libpwc_b200.so.  With this module importable as ``correlation_cuda`` (``sys.modules['correlation_cuda'] = this``, or its
directory first on sys.path under that name) the UNMODIFIED models/correlation_package/correlation.py runs on the B200
kernels: ``correlation.py:4`` imports it, ``:70`` / ``:86`` call forward / backward.

Contract kept: the caller passes EMPTY tensors for rInput1 / rInput2 / output (``input1.new()``); forward
``resize_``s ``output`` to [B, 81, H, W] and writes it in place, returns 1; the padded NHWC scratch tensors are not needed and
are left empty.  dtypes as AT_DISPATCH_FLOATING_TYPES_AND_HALF (correlation_cuda_kernel.cu:386-415): fp32 natively; fp16 / fp64
inputs are converted to fp32 for the kernel -- the reference accumulates in ``float`` for every dtype (.cu:112) -- and the
result is cast back.  Only the PWC-Net configuration is built (pad = md = 4, kernel 1, strides 1, multiply).
No CPU fallback: CUDA tensors only.
"""
import torch

from . import _lib


def _check(input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2, corr_type_multiply):
    if (kernel_size, stride1, stride2, corr_type_multiply) != (1, 1, 1, 1) or pad_size != max_displacement or max_displacement != 4:
        raise NotImplementedError("libpwc_b200 implements the PWC-Net correlation configuration (PWCNet.py:71)")
    if not (input1.is_cuda and input2.is_cuda):
        raise RuntimeError("CUDA call failed")                      # the binding's only error (correlation_cuda.cc:81-83)
    if input1.shape != input2.shape or input1.dim() != 4:
        raise RuntimeError("correlation_cuda: two [B,C,H,W] tensors of the same shape expected")
    if input1.dtype not in (torch.float32, torch.float16, torch.float64):
        raise RuntimeError(f'"correlation_forward" not implemented for {input1.dtype}')


def forward(input1, input2, rInput1, rInput2, output, pad_size, kernel_size, max_displacement, stride1, stride2,
            corr_type_multiply):
    _check(input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2, corr_type_multiply)
    lib = _lib.load()
    B, C, H, W = input1.shape
    a, b = input1.contiguous().float(), input2.contiguous().float()
    with torch.cuda.device_of(input1):
        output.resize_(B, 81, H, W)
        out32 = output if output.dtype == torch.float32 else torch.empty(B, 81, H, W, device=input1.device, dtype=torch.float32)
        e = lib.pwc_corr_fwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out32), B, C, H, W, max_displacement, _lib.stream_handle())
        if e:
            raise RuntimeError("CUDA call failed")
        if out32 is not output:
            output.copy_(out32)
    return 1


def backward(input1, input2, rInput1, rInput2, gradOutput, gradInput1, gradInput2, pad_size, kernel_size, max_displacement,
             stride1, stride2, corr_type_multiply):
    _check(input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2, corr_type_multiply)
    lib = _lib.load()
    B, C, H, W = input1.shape
    a, b, go = input1.contiguous().float(), input2.contiguous().float(), gradOutput.contiguous().float()
    with torch.cuda.device_of(input1):
        gradInput1.resize_(B, C, H, W); gradInput2.resize_(B, C, H, W)
        f32 = gradInput1.dtype == torch.float32
        g1 = gradInput1 if f32 else torch.empty(B, C, H, W, device=input1.device, dtype=torch.float32)
        g2 = gradInput2 if f32 else torch.empty(B, C, H, W, device=input1.device, dtype=torch.float32)
        e = lib.pwc_corr_bwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(go), _lib.ptr(g1), _lib.ptr(g2), B, C, H, W, max_displacement,
                             _lib.stream_handle())
        if e:
            raise RuntimeError("CUDA call failed")
        if not f32:
            gradInput1.copy_(g1); gradInput2.copy_(g2)
    return 1
