"""ctypes binding of the C-ABI library (include/vc_b200.h).

The library is the product: if it is missing, or no sm_100 device is present, calls fail loudly.
There is no CPU or PyTorch fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes
from ctypes import (POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint8, c_uint32,
                    c_void_p)
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "_vc_b200.so"


class VcError(RuntimeError):
    pass


class MatchOptions(ctypes.Structure):
    """vc_match_options (include/vc_b200.h)."""
    _fields_ = [("max_ratio", c_float), ("max_distance", c_float), ("cross_check", c_int32),
                ("inv_norm", c_float), ("rule", c_int32), ("l2_offset", c_float)]


class PairGeometry(ctypes.Structure):
    """vc_pair_geometry (include/vc_b200.h)."""
    _fields_ = [("h", c_float * 9), ("f", c_float * 9), ("mode", c_int32), ("max_h", c_float), ("max_f", c_float)]


GUIDED_SIFTGPU, GUIDED_F, GUIDED_H = 0, 1, 2

# name -> (restype, argtypes); kept in one table so tests can check every exported symbol
SIGNATURES = {
    "vc_last_error": (c_char_p, []),
    "vc_version": (c_char_p, []),
    "vc_launch_count": (c_int64, [c_int]),
    "vc_create": (c_int, [c_int, POINTER(c_void_p)]),
    "vc_destroy": (c_int, [c_void_p]),
    "vc_match_workspace_bytes": (c_size_t, [c_int, c_int]),
    "vc_match_pairs": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                               POINTER(MatchOptions), c_void_p, c_void_p, c_void_p, c_size_t,
                               c_void_p]),
    "vc_match_pairs_top2": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                    POINTER(MatchOptions), c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_size_t, c_void_p]),
    "vc_debug_scores": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                c_void_p, c_int, c_void_p]),
    "vc_match_pairs_host": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int,
                                    c_void_p, c_int, POINTER(MatchOptions), c_void_p, c_void_p]),
    "vc_match_pairs_resident": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                        POINTER(MatchOptions), c_void_p, c_void_p]),
    "vc_match_pairs_guided": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                      c_int, POINTER(MatchOptions), c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "vc_l2_augment": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, POINTER(c_float), c_void_p]),
    "vc_convert_u8_to_bf16": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "vc_convert_f32_to_bf16": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "vc_score_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "vc_score_map": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_int,
                             c_void_p, c_void_p, c_size_t, c_void_p]),
    "vc_score_map_static": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_int,
                                    c_void_p, c_void_p]),
    "vc_select_keypoints": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float,
                                    c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "vc_sample_pack": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64,
                               c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int,
                               POINTER(c_float), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "vc_trainable_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "vc_trainable_select": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_int,
                                    POINTER(c_float), c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_size_t,
                                    c_void_p]),
    "vc_vit_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "vc_vit_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                               c_void_p, c_size_t, c_void_p]),
    "vc_vit_linear": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                              c_int, c_void_p, c_void_p]),
    "vc_vit_linear_ln": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                 c_int, c_void_p, c_void_p, c_void_p]),
    "vc_vit_attention": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "vc_vit_layernorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p,
                                 c_void_p]),
    "vc_head_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "vc_conv_head_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "vc_conv2d": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int,
                          c_void_p, c_void_p, c_void_p, c_void_p]),
    "vc_resize_bilinear": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                   c_void_p]),
    "vc_nchw_to_tokens": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "vc_warp_similarity": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_float,
                                   c_float, c_float, c_void_p, c_void_p]),
    "vc_topk_rows": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p,
                             c_void_p]),
    "vc_hard_negatives": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_float, c_void_p, c_void_p, c_void_p]),
    "vc_sample_bilinear": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p,
                                   c_void_p]),
}


class VitBlock(ctypes.Structure):
    """vc_vit_block (include/vc_b200.h)."""
    _fields_ = [(n, c_void_p) for n in ("ln1_w", "ln1_b", "qkv_w", "qkv_b", "proj_w", "proj_b", "ls1",
                                        "ln2_w", "ln2_b", "fc1_w", "fc1_b", "fc2_w", "fc2_b", "ls2",
                                        "qkv_wf", "qkv_cs", "qkv_bf", "fc1_wf", "fc1_cs", "fc1_bf")]


class VitWeights(ctypes.Structure):
    """vc_vit_weights (include/vc_b200.h)."""
    _fields_ = [("embed_dim", c_int32), ("depth", c_int32), ("num_heads", c_int32), ("patch_kpad", c_int32),
                ("patch_w", c_void_p), ("patch_b", c_void_p), ("cls_token", c_void_p),
                ("pos_embed", c_void_p), ("norm_w", c_void_p), ("norm_b", c_void_p),
                ("blocks", POINTER(VitBlock))]

class HeadWeights(ctypes.Structure):
    """vc_head_weights (include/vc_b200.h)."""
    _fields_ = [("backbone_dim", c_int32), ("descriptor_dim", c_int32),
                ("deconv_w", c_void_p * 2), ("deconv_b", c_void_p * 2),
                ("conv_w", c_void_p * 2), ("conv_scale", c_void_p * 2), ("conv_shift", c_void_p * 2),
                ("trunk_w", c_void_p), ("trunk_scale", c_void_p), ("trunk_shift", c_void_p),
                ("heads_w", c_void_p), ("heads_scale", c_void_p), ("heads_shift", c_void_p),
                ("kp_w", c_void_p), ("kp_b", c_void_p), ("desc_w", c_void_p), ("desc_b", c_void_p)]


DTYPE_U8, DTYPE_F32, DTYPE_BF16 = 0, 1, 2

_lib = None


def load() -> ctypes.CDLL:
    """Load _vc_b200.so (built by vit_colmap_b200.build).  Raises if it is absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise VcError(f"{LIB_PATH} not found: run `python -m vit_colmap_b200.build` "
                          "(there is no fallback implementation)")
        lib = ctypes.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().vc_last_error()
        raise VcError(f"vc_b200 error {rc}: {msg.decode() if msg else '?'}")


def launch_count(reset: bool = False) -> int:
    return int(load().vc_launch_count(1 if reset else 0))
