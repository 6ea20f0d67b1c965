"""ctypes binding of libs2p_sm100.so (include/s2p_b200.h).

The library is built in-tree by ``s2p_b200/csrc/Makefile`` (``__graft_entry__.build()``)
and there is NO fallback: if it is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libs2p_sm100.so")

ACT_NONE, ACT_RELU, ACT_LRELU = 0, 1, 2
IMG_F32_NCHW, IMG_U8_NHWC = 0, 1
RAD_CROP, RAD_TRANSLATE, RAD_CUTOUT, RAD_FLIP, RAD_ROT90 = 0, 1, 2, 3, 4

_vp, _i, _f = C.c_void_p, C.c_int, C.c_float

# name -> (restype, argtypes); mirrors include/s2p_b200.h one to one
SIGNATURES = {
    "s2p_version": (_i, []),
    "s2p_last_error_string": (C.c_char_p, []),
    "s2p_launch_count": (C.c_longlong, []),
    "s2p_conv_stats_rows": (_i, [_i, _i, _i, _i, _i]),
    "s2p_conv_igemm_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _i, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2p_conv_igemm_fwd2": (_i, [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2p_conv_wgrad_workspace_bytes": (C.c_longlong, [_i, _i, _i, _i, _i, _i, _i, _i]),
    "s2p_conv_wgrad": (_i, [_vp, _i, _vp, _i, _vp, _vp, C.c_longlong, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2p_pack_conv_weight": (_i, [_vp, _vp, _i, _i, _i, _i, _f, _i, _vp]),
    "s2p_in_stats_finalize": (_i, [_vp, _i, _vp, _i, _i, _i, _f, _vp]),
    "s2p_mat_norm_fwd": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2p_mat_fused_ct": (_i, [_i]),
    "s2p_mat_prepare": (_i, [_vp, _i, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _f, _vp]),
    "s2p_conv_mat_fused_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2p_mat_norm_bwd_chunks": (_i, [_i, _i, _i, _i]),
    "s2p_mat_norm_bwd": (_i, [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _vp, _i, _vp, _i, _vp, _vp, _vp,
                              _i, _vp, _i, _i, _i, _i, _i, _vp]),
    "s2p_stem_stats_rows": (_i, [_i, _i]),
    "s2p_stem_conv_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp]),
    "s2p_torgb_tanh_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "s2p_torgb_igemm_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "s2p_adam_step": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _f, _f, _f, _f, _i, _vp]),
    "s2p_adam_step_dev": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _f, _f, _f, _f, _vp, _vp]),
    "s2p_pack_conv_weight_dgrad": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "s2p_pack_conv_weights_multi": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "s2p_act_bwd_bias_grad_chunks": (_i, []),
    "s2p_act_bwd_bias_grad": (_i, [_vp, _i, _vp, _i, _i, _vp, _i, _vp, _vp, C.c_longlong, _i, _vp]),
    "s2p_pos_encode_fwd": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "s2p_linear_workspace_floats": (C.c_longlong, [_i, _i, _i]),
    "s2p_linear_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, C.c_longlong, _vp]),
    "s2p_state_encoder_workspace_bytes": (C.c_longlong, [_i, _i]),
    "s2p_state_encoder_static_max_floats": (C.c_longlong, []),
    "s2p_state_encoder_fwd_static": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _i, _vp, _i, _vp]),
    "s2p_state_encoder_fwd": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _i, _vp, _i, _vp, C.c_longlong, _vp]),
    "s2p_adam_step_multi": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _f, _f, _f, _f, _vp, _vp]),
    "s2p_transition_member_floats": (C.c_longlong, [_i, _i, _i]),
    "s2p_ensemble_transition_fwd": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f,
                                         _vp, _vp, _vp, _vp, C.c_longlong, _i, _i, _i, _vp]),
    "s2p_slac_window_gather": (_i, [_vp, _vp, C.c_longlong, _vp, _vp, C.c_longlong, _i, _i, _i, _vp]),
    "s2p_slac_conv1_fwd": (_i, [_vp, _vp, C.c_longlong, _vp, _vp, _vp, _vp, _vp, _i, C.c_longlong, _i, _i, _vp]),
    "s2p_maxpool2_fwd": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp]),
    "s2p_colsum_rows_f32": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "s2p_avgpool3s2_fwd": (_i, [_vp, _vp, C.c_longlong, _i, _i, _vp]),
    "s2p_avgpool3s2_bwd": (_i, [_vp, _vp, C.c_longlong, _i, _i, _vp]),
    "s2p_zero_insert2": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "s2p_tanh_nhwc_to_nchw": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _vp]),
    "s2p_tanh_bwd_nchw_to_nhwc64": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "s2p_nchw_to_nhwc64": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "s2p_nhwc64_to_nchw": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _vp]),
    "s2p_maxpool2_bwd": (_i, [_vp, _i, _vp, _i, _vp, _i, _i, _i, _i, _i, _vp]),
    "s2p_multi_scale_blocks": (_i, []),
    "s2p_multi_scale_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "s2p_multi_scale_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "s2p_l1_mean_chunks": (_i, []),
    "s2p_l1_mean_fwd": (_i, [_vp, _i, _vp, _i, C.c_longlong, _i, _vp, _vp, _vp]),
    "s2p_l1_mean_bwd": (_i, [_vp, _i, _vp, _i, _vp, C.c_longlong, _i, _vp, _i, _vp]),
    "s2p_slac_indices": (_i, [_vp, _vp, _vp, C.c_longlong, _i, C.c_longlong, _vp]),
    "s2p_rad_augment": (_i, [_vp, _vp, _i, C.c_longlong, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "s2p_rad_grayscale": (_i, [_vp, _vp, C.c_longlong, _i, _i, _i, _vp, _vp]),
    "s2p_rad_grayscale_plain": (_i, [_vp, _i, _vp, C.c_longlong, _i, _i, _i, _vp]),
    "s2p_rad_random_convolution": (_i, [_vp, _vp, _vp, C.c_longlong, _i, _i, _i, _vp]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises RuntimeError when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C s2p_b200/csrc).  s2p_b200 has no CPU or PyTorch fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code: int, what: str):
    if code != 0:
        msg = load().s2p_last_error_string()
        raise RuntimeError(f"{what} failed with code {code}: {msg.decode() if msg else '?'}")


def launch_count() -> int:
    return int(load().s2p_launch_count())
