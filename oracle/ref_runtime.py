"""TEST INFRASTRUCTURE -- the reference itself, runnable on the GPU box.

``oracle/build_ref.sh`` (build container) leaves in ``oracle/_ref/`` (git-ignored, ships with the gpurun snapshot):
  * ``correlation_cuda*.so``  the reference's own CUDA correlation (models/correlation_package/correlation_cuda.cc +
    correlation_cuda_kernel.cu with the 7-token F8 patch) compiled for sm_100a;
  * ``refpkg/models/...``     UNMODIFIED copies of models/PWCNet.py and models/correlation_package/correlation.py.
This module imports them.  With a CUDA device the real extension serves ``correlation_cuda.forward`` -- the reference's
hot path exactly as its authors run it (second oracle, GPU baseline).  Without one (the build container, the CPU arm of
bench.py) the SURVEY App. D harness applies: stub module, USE_ONNX_CORRELATION, the /C restatement, identity ``.cuda()``.
Only tests/, __graft_entry__.smoke() and bench.py's baseline legs may import this file.
"""
import glob
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
PKG_DIR = os.path.join(REF_DIR, "refpkg")


def so_path():
    c = glob.glob(os.path.join(REF_DIR, "correlation_cuda*.so"))
    return c[0] if c else None


def available():
    return os.path.isfile(os.path.join(PKG_DIR, "models", "PWCNet.py"))


def cuda_ext_available():
    return so_path() is not None and torch.cuda.is_available()


_loaded = {}


def load(use_cuda_ext=None):
    """Returns (PWCDCNet class, correlation module, using_cuda_ext)."""
    if not available():
        raise RuntimeError("oracle/_ref/refpkg is missing: run oracle/build_ref.sh in the build container")
    if use_cuda_ext is None:
        use_cuda_ext = cuda_ext_available()
    if _loaded:
        if _loaded["ext"] != use_cuda_ext:
            raise RuntimeError("reference already loaded in the other mode in this process")
        return _loaded["cls"], _loaded["corr"], _loaded["ext"]
    sys.dont_write_bytecode = True
    if use_cuda_ext:
        if REF_DIR not in sys.path:
            sys.path.insert(0, REF_DIR)
        import correlation_cuda  # noqa: F401  (the reference's own extension, built for sm_100a)
    else:
        sys.modules.setdefault("correlation_cuda", types.ModuleType("correlation_cuda"))
        if not torch.cuda.is_available():
            torch.Tensor.cuda = lambda self, *a, **k: self
    if PKG_DIR not in sys.path:
        sys.path.insert(0, PKG_DIR)
    for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
        del sys.modules[k]
    from models.PWCNet import PWCDCNet  # noqa: E402
    from models.correlation_package import correlation as corr_mod  # noqa: E402
    if not getattr(corr_mod, "_oracle_patched", False):
        _sum = corr_mod.correlation_onnx                 # F1: the CUDA kernel divides by C, the torch path does not
        corr_mod.correlation_onnx = lambda a, b, *k: _sum(a, b, *k) / a.shape[1]
        corr_mod._oracle_patched = True
    corr_mod.USE_ONNX_CORRELATION = not use_cuda_ext
    _loaded.update(cls=PWCDCNet, corr=corr_mod, ext=use_cuda_ext)
    return PWCDCNet, corr_mod, use_cuda_ext


class cpu_execution:
    """Run the loaded reference on CPU tensors inside a process that also has CUDA (bench.py's CPU baseline on the GPU
    box): the torch shift-and-dot correlation (/C) instead of the extension, and PWCNet.py:167's ``mask.cuda()`` as identity."""

    def __enter__(self):
        _, corr_mod, _ = load()
        self.corr_mod, self.flag, self.cuda = corr_mod, corr_mod.USE_ONNX_CORRELATION, torch.Tensor.cuda
        corr_mod.USE_ONNX_CORRELATION = True
        torch.Tensor.cuda = lambda t, *a, **k: t
        return self

    def __exit__(self, *exc):
        self.corr_mod.USE_ONNX_CORRELATION = self.flag
        torch.Tensor.cuda = self.cuda
        return False


def reference_model(seed=0, device=None, use_cuda_ext=None):
    """The reference's PWCDCNet with its own init under torch.manual_seed(seed), in eval mode."""
    cls, _, ext = load(use_cuda_ext)
    torch.manual_seed(seed)
    m = cls().eval()
    if device is not None:
        m = m.to(device)
    return m


def exact_fp32():
    """The reference under its own defaults lets cuDNN use TF32 for convolutions on Ampere+; parity runs switch it off."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
