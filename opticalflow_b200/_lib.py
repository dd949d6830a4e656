"""ctypes binding of libpwc_b200.so (the C ABI declared in include/pwc_b200.h).

The product path has no CPU fallback: if the library is missing or a CUDA device is absent the
callers raise.  ``load()`` builds the library in-tree when the sources are newer (nvcc, sm_100a).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpwc_b200.so")
_lib = None

F32, BF16 = 0, 1
NHWC, NCHW = 0, 1

_vp, _i, _l, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float

SIGNATURES = {
    "pwc_b200_arch": (_i, []),
    "pwc_b200_version": (ctypes.c_char_p, []),
    "pwc_corr_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "pwc_corr_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "pwc_warp_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _f, _vp]),
    "pwc_warp_corr_lrelu_fwd": (_i, [_vp, _l, _vp, _l, _vp, _l, _f, _i, _f, _f, _vp, _l,
                                     _i, _i, _i, _i, _i, _i, _i, _vp]),
    "pwc_conv3x3_f32_fwd": (_i, [_vp, _l, _i, _vp, _l, _i, _i, _vp, _vp, _vp, _l, _i, _vp, _l, _i, _i,
                                 _i, _i, _i, _i, _i, _i, _f, _vp, _l, _vp]),
    "pwc_pack_conv_weight_f32": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "pwc_deconv4x4s2_c2_f32_fwd": (_i, [_vp, _l, _i, _i, _vp, _vp, _vp, _l, _i, _vp, _l, _i, _i, _i, _vp]),
    "pwc_pack_deconv_weight_f32": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "pwc_tc_conv_create": (_i, [ctypes.POINTER(_vp), _vp, _l, _i, _vp, _l, _i, _i, _i, _i, _vp, _vp, _i, _i,
                                _vp, _l, _i, _vp, _l, _i, _i, _vp, _l, _i, _i, _i, _i, _i, _f]),
    "pwc_tc_conv_create_split": (_i, [ctypes.POINTER(_vp), _vp, _l, _i, _vp, _l, _i, _i, _i, _i, _vp, _vp, _i, _i,
                                      _vp, _l, _i, _vp, _l, _i, _i, _vp, _l, _i, _i, _i, _i, _i, _f, _i, _i, _i, _i]),
    "pwc_conv3x3_thin_split_fwd": (_i, [_vp, _l, _i, _i, _vp, _vp, _i, _vp, _l, _i, _i, _i, _i, _f, _vp]),
    "pwc_conv3x3_thin_split_strided_fwd": (_i, [_vp, _l, _i, _i, _vp, _vp, _i, _vp, _l, _i, _i, _i, _i, _i, _f, _vp]),
    "pwc_conv1a_split_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _vp]),
    "pwc_deconv4x4s2_c2_split_fwd": (_i, [_vp, _i, _l, _i, _i, _vp, _vp, _vp, _i, _l, _i, _vp, _l, _i, _i, _i, _i, _i, _vp]),
    "pwc_copy_channels_h16": (_i, [_vp, _l, _i, _vp, _l, _i, _i, _l, _vp]),
    "pwc_tc_conv_launch": (_i, [_vp, _vp]),
    "pwc_tc_conv_launch_static_splitk": (_i, [_vp, _i, _vp]),
    "pwc_tc_conv_set_skip": (_i, [_vp, _vp]),
    "pwc_tc_conv_set_kk_skip": (_i, [_vp, _vp]),
    "pwc_tc_conv_set_out_groups": (_i, [_vp, _i, _l]),
    "pwc_tc_conv_destroy": (None, [_vp]),
    "pwc_tc_conv_num_launches": (_i, [_vp]),
    "pwc_tc_conv_set_head_rows": (_i, [_vp, _vp, _i]),
    "pwc_tc_conv_set_partial": (_i, [_vp, _vp, _i]),
    "pwc_set_half_format": (_i, [_i]),
    "pwc_get_half_format": (_i, []),
    "pwc_debug_halo_trace": (_i, [_vp, _i]),
    "pwc_debug_corr_trace": (_i, [_vp, _i]),
    "pwc_debug_wm_trace": (_i, [_vp, _i]),
    "pwc_pack_conv_weight_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "pwc_conv1a_bf16_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    "pwc_deconv4x4s2_c2_fwd": (_i, [_vp, _i, _l, _i, _i, _vp, _vp, _vp, _i, _l, _i, _vp, _l, _i, _i, _i, _vp]),
    "pwc_pack_input_f32": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "pwc_unpack_flow_f32": (_i, [_vp, _l, _i, _vp, _i, _i, _i, _f, _vp]),
    "pwc_conv3x3_mma_fwd": (_i, [_vp, _l, _i, _vp, _l, _i, _i, _i, _vp, _vp, _i, _i, _vp, _l, _i, _vp, _l, _i, _i,
                                 _i, _i, _i, _i, _f, _vp]),
    "pwc_conv3x3_thin_fwd": (_i, [_vp, _l, _i, _vp, _l, _i, _i, _i, _vp, _vp, _i, _vp, _l, _i, _vp, _l, _i, _i,
                                  _i, _i, _i, _i, _f, _vp]),
    "pwc_conv3x3_flow_head_fwd": (_i, [_vp, _l, _i, _i, _vp, _vp, _vp, _l, _vp, _i, _i, _i, _vp]),
    "pwc_preprocess_u8_fwd": (_i, [_vp, _vp, _l, _l, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "pwc_preprocess_u8_norm_fwd": (_i, [_vp, _vp, _l, _l, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "pwc_preprocess_frames_u8_fwd": (_i, [_vp, _l, _l, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "pwc_conv1a_frames_bf16_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    "pwc_flow_postprocess_fwd": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _i, _f, _i, _vp]),
    "pwc_flow_metrics_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _vp, _vp]),
}


class PwcLibraryError(RuntimeError):
    pass


def load(build_if_needed=True):
    """Load (building first if sources are newer and nvcc exists).  Raises PwcLibraryError if unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_needed:
        try:
            from . import build as _build
            if _build.needs_build() and os.path.exists(_build.NVCC):
                _build.build()
        except Exception as e:  # noqa: BLE001
            if not os.path.isfile(LIB_PATH):
                raise PwcLibraryError(f"libpwc_b200.so is not built and could not be built: {e}") from e
    if not os.path.isfile(LIB_PATH):
        raise PwcLibraryError(f"{LIB_PATH} not found: run `python -m opticalflow_b200.build`")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise PwcLibraryError(f"libpwc_b200.so does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(err, what):
    if err != 0:
        raise RuntimeError(f"{what}: CUDA call failed (cudaError {err})")


def ptr(t):
    """data pointer of a torch tensor (or None)."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_handle():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
