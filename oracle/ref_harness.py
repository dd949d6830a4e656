"""Import the UNMODIFIED reference PWCDCNet from /root/reference on a CPU-only box.

Build-container only (``/root/reference`` does not exist on the GPU box): used by
``oracle/make_golden.py`` to pin ``oracle/pwcnet_oracle.py`` and to write ``tests/golden``.
Recipe: SURVEY.md Appendix D.  No reference file is modified.
  * correlation.py:4 imports the compiled ``correlation_cuda`` unconditionally -> stub module;
  * ``USE_ONNX_CORRELATION = True`` selects the pure-torch shift-and-dot (correlation.py:105-110);
  * that path sums over channels, the CUDA kernel divides by C (correlation_cuda_kernel.cu:104,143)
    -> wrap it with ``/ C``;
  * PWCNet.py:167 calls ``.cuda()`` on the mask -> identity shim on a CPU box.
"""
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("PWC_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "models", "PWCNet.py"))


def load_reference():
    """Returns (PWCDCNet class, correlation module)."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    sys.dont_write_bytecode = True
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    sys.modules.setdefault("correlation_cuda", types.ModuleType("correlation_cuda"))
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    from models.PWCNet import PWCDCNet  # noqa: E402
    from models.correlation_package import correlation as corr_mod  # noqa: E402
    if not getattr(corr_mod, "_oracle_patched", False):
        corr_mod.USE_ONNX_CORRELATION = True
        _sum = corr_mod.correlation_onnx
        corr_mod.correlation_onnx = lambda a, b, *k: _sum(a, b, *k) / a.shape[1]
        corr_mod._oracle_patched = True
    return PWCDCNet, corr_mod


def reference_model(seed=0):
    PWCDCNet, _ = load_reference()
    torch.manual_seed(seed)
    return PWCDCNet().eval()
